"""CPU tests of the drop-in boundary and the host logic (no compute calls):
  * libbandcu.so / libband_c.so load and export every symbol include/*.h declares;
  * builder / validation status codes follow the reference (image_processor.cc:25-67,
    image_operator.cc validation) -- checked against the oracle's restatement of the same rules;
  * without a CUDA device the product FAILS LOUDLY (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT


def _declared(header, prefix):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    # only exported declarations: `BANDCU_API <ret> name(` / `BAND_CAPI_EXPORT extern <ret> name(`
    names = set(re.findall(r"(?:BANDCU_API|BAND_CAPI_EXPORT)[^;(]*?\b(%s\w+)\s*\(" % prefix, text))
    return names


def test_bandcu_exports_every_declared_symbol():
    from band_b200 import cuda_shim as S
    L = S.lib()
    declared = _declared("bandcu.h", "bandcu_")
    assert declared, "no declarations parsed"
    for name in sorted(declared):
        assert hasattr(L, name), name
    assert declared == set(S.EXPORTED)


def test_band_c_exports_every_declared_symbol():
    from band_b200 import band_c as B
    L = B.lib()
    declared = _declared("band_c_api.h", "Band")
    for name in sorted(declared):
        assert hasattr(L, name), name
    assert declared == set(B.EXPORTED)


def test_reference_c_api_names_present():
    """The exported names a dlopen() user of the reference binds (band/c/c_api_buffer.h:103-121)."""
    from band_b200 import band_c as B
    L = B.lib()
    for name in ("BandBufferCreate", "BandBufferDelete", "BandBufferSetFromRawData", "BandBufferSetFromYUVData",
                 "BandImageProcessorBuilderCreate", "BandImageProcessorBuilderDelete", "BandImageProcessorBuilderBuild",
                 "BandAddOperator", "BandImageProcessorProcess", "BandImageProcessorDelete", "BandTensorGetData",
                 "BandTensorGetDims", "BandTensorGetNumDims", "BandTensorGetBytes", "BandTensorGetType"):
        assert hasattr(L, name), name


def _testing():
    path = os.path.join(ROOT, "band_b200", "lib", "libband_testing.so")
    T = C.CDLL(path)
    return T


class _OpDesc(C.Structure):
    _fields_ = [("kind", C.c_int), ("i", C.c_int * 4), ("f", C.c_float * 2), ("inplace", C.c_int)]


def _descs(ops):
    arr = (_OpDesc * max(1, len(ops)))()
    for k, op in enumerate(ops):
        arr[k].kind = op[0]
        for j, v in enumerate(list(op[1:5])):
            arr[k].i[j] = int(v)
        for j, v in enumerate(list(op[5:7])):
            arr[k].f[j] = float(v)
    return arr


@pytest.mark.parametrize("ops,ok", [
    ([(1, 224, 224), (5, 0, 0, 0, 0, 127.5, 127.5)], True),
    ([(1, 224, 224), (1, 100, 100)], False),          # duplicate operator type (image_processor.cc:31-43)
    ([(0, 0, 0, 9, 9), (2, 90), (0, 1, 1, 5, 5)], False),
    ([], True),                                        # empty -> AutoConvert (image_processor.cc:62-64)
    ([(2, 90), (0, 0, 0, 9, 9), (4, 1), (1, 32, 32), (5, 0, 0, 0, 0, 0.0, 1.0)], True),
])
def test_builder_status_matches_oracle(O, ops, ok):
    T = _testing()
    T.bandt_build_status.argtypes = [C.POINTER(_OpDesc), C.c_int]
    got = T.bandt_build_status(_descs(ops), len(ops))
    st = C.c_int(0)
    p = O.lib().orc_processor_build(O.make_descs(ops), len(ops), C.byref(st))
    if p:
        O.lib().orc_processor_free(p)
    assert (got == 0) == ok
    assert got == st.value


def test_add_operator_argument_counts():
    """c_api_buffer.cc:107-196: a wrong `count` for the field is kBandErr."""
    from band_b200 import band_c as B
    L = B.lib()
    b = L.BandImageProcessorBuilderCreate()
    try:
        assert B.add_operator(b, B.BAND_RESIZE, 224) == B.kBandErr
        assert B.add_operator(b, B.BAND_CROP, 0, 0, 1) == B.kBandErr
        assert B.add_operator(b, B.BAND_ROTATE, 90, 1) == B.kBandErr
        assert B.add_operator(b, B.BAND_NORMALIZE, 1.0) == B.kBandErr
        assert B.add_operator(b, B.BAND_RESIZE, 224, 224) == B.kBandOk
        assert B.add_operator(b, B.BAND_NORMALIZE, 127.5, 127.5) == B.kBandOk
        assert B.add_operator(b, B.BAND_DATA_TYPE_CONVERT) == B.kBandOk
        assert B.add_operator(None, B.BAND_RESIZE, 1, 1) == B.kBandErr
    finally:
        L.BandImageProcessorBuilderDelete(b)


def test_null_handles_are_errors():
    from band_b200 import band_c as B
    L = B.lib()
    assert L.BandImageProcessorProcess(None, None, None) == B.kBandErr
    assert L.BandBufferSetFromRawData(None, None, 1, 1, B.kBandRGB) == B.kBandErr
    assert not L.BandImageProcessorBuilderBuild(None)


def _no_gpu():
    import torch
    return not torch.cuda.is_available()


@pytest.mark.skipif(not _no_gpu(), reason="checks the behaviour WITHOUT a device")
def test_no_cpu_fallback_without_device():
    """No CUDA device: plan creation / processing must fail with an error, never compute on the CPU."""
    from band_b200 import band_c as B
    from band_b200 import cuda_shim as S
    n = C.c_int(-1)
    rc = S.lib().bandcu_device_count(C.byref(n))
    assert rc != 0 or n.value == 0
    frame = np.zeros(B.frame_bytes(64, 48, B.kBandNV12), dtype=np.uint8)
    ops = [(B.BAND_COLOR_SPACE_CONVERT, B.kBandRGB), (B.BAND_RESIZE, 22, 22)]
    try:
        st, out = B.run_chain(ops, frame[None], 64, 48, B.kBandNV12, np.uint8, (1, 22, 22, 3))
    except B.BandError:
        return  # tensor / processor creation refused: loud failure
    assert st != B.kBandOk and out is None


def test_no_alloc_scratch_size_is_the_headers():
    """The Python shim's constant and include/bandcu.h agree on the static scratch size of bandcu_two_stage_run_noalloc."""
    import re
    from band_b200 import cuda_shim as S
    text = open(os.path.join(ROOT, "include", "bandcu.h")).read()
    m = re.search(r"#define\s+BANDCU_NOALLOC_SCRATCH_BYTES\s+\((\d+)u\s*<<\s*(\d+)\)", text)
    assert m, "BANDCU_NOALLOC_SCRATCH_BYTES not found in include/bandcu.h"
    assert int(m.group(1)) << int(m.group(2)) == S.NOALLOC_SCRATCH_BYTES == 8 << 20
