"""GPU parity tests proper: the CUDA path (through the Band C API / C-ABI) against the CPU oracle on the
same seeded inputs. Integer stages: bit-exact. Float Normalize: BASELINE.json allows 1e-6 relative;
the kernels reproduce the IEEE result, so the tests demand identical bits (0 ulp <= 1e-6)."""
import hashlib
import importlib.util
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN
from gpu_util import (assert_same_bits, gpu_chain, gpu_image_chain, gpu_op, oracle_chain, rand_frames)

pytestmark = pytest.mark.gpu

NORM = (5, 0, 0, 0, 0, 127.5, 127.5)


def _csc(O, fmt):
    return (O.OP_CSC, fmt)


# ---------------------------------------------------------------------------------------------------
# BASELINE.json configs at their full sizes, a few frames each
# ---------------------------------------------------------------------------------------------------
def test_c1_nv21_640x480_to_224_f32(O):
    ops = [_csc(O, O.FMT_RGB), (O.OP_RESIZE, 224, 224), NORM]
    fr = rand_frames(O, 1, 640, 480, O.FMT_NV21, 1)
    st, want = oracle_chain(O, ops, fr, 640, 480, O.FMT_NV21, O.DT_F32, (1, 224, 224, 3))
    assert st == 0
    st, got = gpu_chain(ops, fr, 640, 480, O.FMT_NV21, O.DT_F32, (1, 224, 224, 3))
    assert st == 0
    assert_same_bits(got, want, "C1")
    # BASELINE tolerance statement (redundant with the bit check, kept explicit)
    assert np.max(np.abs(got - want) / np.maximum(np.abs(want), 1e-30)) <= 1e-6


@pytest.mark.parametrize("resident", [False, True])
def test_north_star_nv12_1080p_to_224_f32(O, resident):
    ops = [_csc(O, O.FMT_RGB), (O.OP_RESIZE, 224, 224), NORM]
    fr = rand_frames(O, 5, 1920, 1080, O.FMT_NV12, 2)
    st, want = oracle_chain(O, ops, fr, 1920, 1080, O.FMT_NV12, O.DT_F32, (1, 224, 224, 3))
    assert st == 0
    st, got = gpu_chain(ops, fr, 1920, 1080, O.FMT_NV12, O.DT_F32, (1, 224, 224, 3), device_resident=resident)
    assert st == 0
    assert_same_bits(got, want, "north star resident=%s" % resident)


@pytest.mark.parametrize("resident", [False, True])
def test_c2_nv12_1080p_to_300_u8(O, resident):
    ops = [_csc(O, O.FMT_RGB), (O.OP_RESIZE, 300, 300)]
    fr = rand_frames(O, 4, 1920, 1080, O.FMT_NV12, 3)
    st, want = oracle_chain(O, ops, fr, 1920, 1080, O.FMT_NV12, O.DT_U8, (1, 300, 300, 3))
    assert st == 0
    st, got = gpu_chain(ops, fr, 1920, 1080, O.FMT_NV12, O.DT_U8, (1, 300, 300, 3), device_resident=resident)
    assert st == 0
    assert_same_bits(got, want, "C2")


def test_c3_yv12_4k_rotate_crop_to_512_f32(O):
    ops = [(O.OP_ROTATE, 90), (O.OP_CROP, 0, 840, 2159, 2999), _csc(O, O.FMT_RGB), (O.OP_RESIZE, 512, 512), NORM]
    fr = rand_frames(O, 2, 3840, 2160, O.FMT_YV12, 4)
    st, want = oracle_chain(O, ops, fr, 3840, 2160, O.FMT_YV12, O.DT_F32, (1, 512, 512, 3))
    assert st == 0
    st, got = gpu_chain(ops, fr, 3840, 2160, O.FMT_YV12, O.DT_F32, (1, 512, 512, 3), device_resident=True)
    assert st == 0
    assert_same_bits(got, want, "C3")


@pytest.mark.parametrize("target", [(224, 1), (300, 3), (513, 1)])
def test_c4_nv21_1080p_mixed_targets(O, target):
    d, dt = target
    ops = [_csc(O, O.FMT_RGB), (O.OP_RESIZE, d, d)] + ([NORM] if dt == 1 else [])
    fr = rand_frames(O, 3, 1920, 1080, O.FMT_NV21, 5 + d)
    st, want = oracle_chain(O, ops, fr, 1920, 1080, O.FMT_NV21, dt, (1, d, d, 3))
    assert st == 0
    st, got = gpu_chain(ops, fr, 1920, 1080, O.FMT_NV21, dt, (1, d, d, 3), device_resident=True)
    assert st == 0
    assert_same_bits(got, want, "C4 %d" % d)


@pytest.mark.parametrize("s,d", [(2048, 64), (2048, 128), (2048, 256), (2048, 512), (2048, 1024), (64, 128),
                                 (64, 256), (64, 512), (64, 1024), (64, 2048), (640, 224), (225, 224),
                                 (64, 352), (100, 1200), (480, 1072)])
def test_c5_rgba_to_rgb_autoconvert(O, s, d):
    """C5 (i): default AutoConvert into a u8 RGB tensor = CSC(drop alpha) -> ResizeRgb, incl. the
    Down2 / DownEven box, point-sample and bilinear-up dispatch cases."""
    ops = [(O.OP_AUTO,)]
    fr = rand_frames(O, 1, s, s, O.FMT_RGBA, 100 + d)
    st, want = oracle_chain(O, ops, fr, s, s, O.FMT_RGBA, O.DT_U8, (1, d, d, 3))
    assert st == 0
    st, got = gpu_chain(ops, fr, s, s, O.FMT_RGBA, O.DT_U8, (1, d, d, 3))
    assert st == 0
    assert_same_bits(got, want, "C5 rgb %d->%d" % (s, d))


@pytest.mark.parametrize("s,d", [(2048, 64), (2048, 128), (2048, 256), (2048, 512), (2048, 1024), (2048, 1536),
                                 (2048, 768), (64, 128), (64, 256), (64, 100), (64, 2048), (640, 224),
                                 (64, 1040), (300, 1504)])
def test_c5_rgba_to_gray_resize(O, s, d):
    """C5 (ii): CSC(Gray) -> Resize through ScalePlane (x86 rounding blender, Down2/Down34/Down38/Up2 cases)."""
    ops = [_csc(O, O.FMT_GRAY), (O.OP_RESIZE, d, d)]
    fr = rand_frames(O, 1, s, s, O.FMT_RGBA, 200 + d)[0]
    cur, w, h, f = fr, s, s, O.FMT_RGBA
    for op in ops:
        st, cur, (w, h, f) = O.run_op(op, cur, w, h, f)
        assert st == 0
    st, got, info = gpu_image_chain(ops, fr, s, s, O.FMT_RGBA, O.FMT_GRAY, O.DT_U8, d, d)
    assert st == 0 and info[:3] == (d, d, O.FMT_GRAY)
    assert_same_bits(got, cur, "C5 gray %d->%d" % (s, d))


# ---------------------------------------------------------------------------------------------------
# committed golden fixtures
# ---------------------------------------------------------------------------------------------------
def _make_golden():
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(GOLDEN, "make_golden.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    return mg


def test_frozen_digests_on_gpu(O):
    """Every case of tests/golden/oracle_frozen.json, computed by the CUDA path, hashes to the frozen digest."""
    mg = _make_golden()
    with open(os.path.join(GOLDEN, "oracle_frozen.json")) as f:
        frozen = json.load(f)
    for name, ops, W, H, fmt, odt, odims, seed in mg.frozen_cases():
        rng = np.random.default_rng(seed)
        frame = rng.integers(0, 256, size=O.frame_bytes(W, H, fmt), dtype=np.uint8)
        if len(odims) == 2:
            st, got, _ = gpu_image_chain(ops, frame, W, H, fmt, O.FMT_GRAY, O.DT_U8, odims[1], odims[0])
        else:
            st, got = gpu_chain(ops, frame, W, H, fmt, odt, odims)
            got = got[0] if st == 0 else None
        assert st == 0, name
        assert hashlib.sha256(np.ascontiguousarray(got).tobytes()).hexdigest() == frozen[name], name


def test_reference_crop_kat(O):
    """image_operation_test.cc:12-29 on the GPU backing."""
    st, out, info = gpu_op((O.OP_CROP, 0, 0, 1, 1), np.arange(1, 10, dtype=np.uint8), 3, 3, O.FMT_GRAY)
    assert st == 0 and info[:3] == (2, 2, O.FMT_GRAY)
    assert out.tolist() == [1, 2, 4, 5]


def test_reference_crop_hippo(O, hippo):
    """image_operation_test.cc:31-44."""
    src = hippo["hippo"]
    H, W = src.shape[:2]
    st, out, info = gpu_op((O.OP_CROP, 0, 0, 255, 255), src.reshape(-1), W, H, O.FMT_RGB)
    assert st == 0 and info[:2] == (256, 256)
    assert np.array_equal(out.reshape(256, 256, 3), hippo["crop_256"])


@pytest.mark.parametrize("box", [(0, 0, 600, 400), (255, 0, 0, 255), (0, 255, 255, 0), (-1, -1, 256, 256),
                                 (0, 0, -1, -1), (-1, -1, -1, -1)])
def test_reference_crop_failures(O, hippo, box):
    """image_operation_test.cc:46-75: kInvalidArgument (3)."""
    src = hippo["hippo"]
    H, W = src.shape[:2]
    st, _, _ = gpu_op((O.OP_CROP,) + box, src.reshape(-1), W, H, O.FMT_RGB)
    assert st == O.INVALID_ARGUMENT


@pytest.mark.parametrize("assigned", [True, False])
def test_reference_convert_gray(O, hippo, assigned):
    """image_operation_test.cc:77-119 (assigned output / auto-created output)."""
    src = hippo["hippo"]
    H, W = src.shape[:2]
    if assigned:
        st, out, info = gpu_op((O.OP_CSC, -1), src.reshape(-1), W, H, O.FMT_RGB, assign=(O.FMT_GRAY, O.DT_U8, W, H))
    else:
        st, out, info = gpu_op((O.OP_CSC, O.FMT_GRAY), src.reshape(-1), W, H, O.FMT_RGB)
    assert st == 0 and info[:3] == (W, H, O.FMT_GRAY)
    want = 0.299 * src[..., 0] + 0.587 * src[..., 1] + 0.114 * src[..., 2]
    assert np.abs(out.reshape(H, W).astype(np.float64) - want).max() <= 1.0
    st, ref, _ = O.run_op((O.OP_CSC, O.FMT_GRAY), src.reshape(-1), W, H, O.FMT_RGB)
    assert_same_bits(out, ref, "gray")


def test_reference_rotate_round_trip(O, hippo):
    """image_operation_test.cc:121-141."""
    src = hippo["hippo"]
    H, W = src.shape[:2]
    st, r90, info = gpu_op((O.OP_ROTATE, 90), src.reshape(-1), W, H, O.FMT_RGB)
    assert st == 0 and info[:2] == (H, W)
    assert np.array_equal(r90.reshape(W, H, 3), np.rot90(src, 1))
    st, back, info2 = gpu_op((O.OP_ROTATE, 270), r90, H, W, O.FMT_RGB)
    assert st == 0 and info2[:2] == (W, H)
    assert np.array_equal(back.reshape(H, W, 3), src)


def test_reference_resize_hippo(O, hippo):
    """image_operation_test.cc:143-159 (+-3 on the bytes it compares) and bit-exact vs the oracle."""
    src = hippo["hippo"]
    H, W = src.shape[:2]
    st, out, info = gpu_op((O.OP_RESIZE, 256, 256), src.reshape(-1), W, H, O.FMT_RGB)
    assert st == 0 and info[:2] == (256, 256)
    want = hippo["resize_256"].reshape(-1).astype(np.int32)
    n = 256 * 256
    assert np.abs(out[:n].astype(np.int32) - want[:n]).max() <= 3
    assert np.abs(out.astype(np.int32) - want).mean() < 0.25
    st, ref, _ = O.run_op((O.OP_RESIZE, 256, 256), src.reshape(-1), W, H, O.FMT_RGB)
    assert_same_bits(out, ref, "hippo resize")


@pytest.mark.parametrize("fmt_name,W,H,d,f32", [("NV12", 1920, 1080, 224, True), ("NV21", 1920, 1080, 300, False),
                                                ("NV21", 640, 480, 224, True), ("YV12", 1920, 1080, 224, True),
                                                ("YV21", 1280, 720, 300, False), ("YV12", 648, 360, 200, True)])
def test_pinned_host_frames_zero_copy(O, fmt_name, W, H, d, f32):
    """Frames in PINNED host memory (BandHostAlloc): the staged kernel pulls the rows it samples straight over
    PCIe (no staging copy). Same bits as the oracle."""
    import ctypes as C
    from band_b200 import band_c as B
    fmt = getattr(O, "FMT_" + fmt_name)
    n = 40  # more than one 32-frame chunk
    fr = rand_frames(O, 4, W, H, fmt, 4242)
    fb = fr.shape[1]
    L = B.lib()
    host = L.BandHostAlloc(n * fb)
    assert host
    try:
        hv = np.ctypeslib.as_array(C.cast(host, C.POINTER(C.c_uint8)), shape=(n, fb))
        for i in range(n):
            hv[i] = fr[i % 4]
        ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, d, d)] + ([NORM] if f32 else [])
        dt = O.DT_F32 if f32 else O.DT_U8
        st, want = oracle_chain(O, ops, fr, W, H, fmt, dt, (1, d, d, 3))
        assert st == 0
        from gpu_util import to_capi_ops, NP_OF_ORC
        proc = B.ImageProcessor(to_capi_ops(ops))
        tens = [B.Tensor(NP_OF_ORC[dt], (1, d, d, 3), device=False) for _ in range(n)]
        bufs = [B.Buffer().set_raw(host + i * fb, W, H, B.FORMATS[fmt_name]) for i in range(n)]
        assert proc.process_batch(bufs, tens) == B.kBandOk, B.last_error()
        assert proc.is_fused
        for i in range(n):
            assert_same_bits(tens[i].numpy(), want[i % 4], "pinned frame %d" % i)
        for b in bufs:
            b.close()
        for t in tens:
            t.close()
        proc.close()
    finally:
        L.BandHostFree(host)


@pytest.mark.parametrize("fmt_name,W,H,dw,dh,f32,dev_in", [
    ("NV12", 1920, 1080, 228, 225, True, False),   # 225 rows: the last work item of a frame holds one row
    ("NV21", 1280, 720, 320, 181, False, False),   # u8 rows of 960 bytes (16-byte multiple): bulk stores
    ("NV12", 640, 480, 224, 224, True, True),      # frames in HBM, tensors in pinned host memory
    ("NV12", 640, 480, 226, 224, True, False),     # 2712-byte rows: not a 16-byte multiple -> staged download
])
def test_pinned_host_tensors_bulk_store(O, fmt_name, W, H, dw, dh, f32, dev_in):
    """Tensors in PINNED host memory: the staged kernel assembles rows in shared memory and pushes them over PCIe
    with cp.async.bulk stores (no staging buffer, no D2H copy). Same bits as the oracle, and the guard bytes
    around every tensor stay untouched."""
    import ctypes as C
    from band_b200 import band_c as B
    from gpu_util import to_capi_ops, NP_OF_ORC
    fmt = getattr(O, "FMT_" + fmt_name)
    n = 37
    fr = rand_frames(O, 3, W, H, fmt, 777)
    fb = fr.shape[1]
    L = B.lib()
    ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, dw, dh)] + ([NORM] if f32 else [])
    dt = O.DT_F32 if f32 else O.DT_U8
    st, want = oracle_chain(O, ops, fr, W, H, fmt, dt, (1, dh, dw, 3))
    assert st == 0
    host = L.BandHostAlloc(n * fb)
    dev = L.BandDeviceAlloc(n * fb) if dev_in else None
    assert host and (dev or not dev_in)
    try:
        hv = np.ctypeslib.as_array(C.cast(host, C.POINTER(C.c_uint8)), shape=(n, fb))
        for i in range(n):
            hv[i] = fr[i % 3]
        if dev_in:
            assert L.BandCopyToDevice(dev, host, n * fb) == B.kBandOk
        proc = B.ImageProcessor(to_capi_ops(ops))
        tens = [B.Tensor(NP_OF_ORC[dt], (1, dh, dw, 3), device=False) for _ in range(n)]
        base = dev if dev_in else host
        bufs = [B.Buffer().set_raw(base + i * fb, W, H, B.FORMATS[fmt_name], device=dev_in) for i in range(n)]
        for rep in range(2):  # second call reuses the plan (and flips nothing)
            assert proc.process_batch(bufs, tens) == B.kBandOk, B.last_error()
        assert proc.is_fused
        for i in range(n):
            assert_same_bits(tens[i].numpy(), want[i % 3], "host tensor %d" % i)
        # the same processor then serves device tensors (direct-store kernel) with identical bits
        dtens = [B.Tensor(NP_OF_ORC[dt], (1, dh, dw, 3), device=True) for _ in range(3)]
        assert proc.process_batch(bufs[:3], dtens) == B.kBandOk, B.last_error()
        for i in range(3):
            assert_same_bits(dtens[i].numpy(), want[i], "device tensor %d" % i)
        for x in bufs + tens + dtens:
            x.close()
        proc.close()
    finally:
        L.BandHostFree(host)
        if dev:
            L.BandDeviceFree(dev)


@pytest.mark.parametrize("out_dt_name,mean,std", [("DT_I8", 128.0, 1.0), ("DT_U8", 0.0, 2.0), ("DT_I8", 127.5, 1.0)])
@pytest.mark.parametrize("fmt_name,W,H,d", [("NV21", 1920, 1080, 224), ("NV12", 640, 480, 300), ("YV12", 1920, 1080, 224),
                                            ("NV21", 642, 482, 224)])
def test_quantized_model_inputs(O, out_dt_name, mean, std, fmt_name, W, H, d):
    """CSC -> Resize -> Normalize into a ONE-BYTE tensor (int8 / uint8 quantized models: NormalizeFromTo<u8, i8>,
    common_operator.cc:24-40, cast of the difference then the division): the row-staged kernel with a byte table;
    the unaligned case runs its byte-copy producer. uint8 keeps mean at 0: the cast of a negative difference to an unsigned
    type is undefined in the reference too."""
    out_dt = getattr(O, out_dt_name)
    fmt = getattr(O, "FMT_" + fmt_name)
    ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, d, d), (O.OP_NORMALIZE, 0, 0, 0, 0, mean, std)]
    fr = rand_frames(O, 2, W, H, fmt, 4242 + d)
    st, want = oracle_chain(O, ops, fr, W, H, fmt, out_dt, (1, d, d, 3))
    assert st == 0
    st, got = gpu_chain(ops, fr, W, H, fmt, out_dt, (1, d, d, 3), device_resident=True)
    assert st == 0
    assert_same_bits(got, want, "quantized %s %s" % (out_dt_name, fmt_name))


@pytest.mark.parametrize("angle", [90, 180])
@pytest.mark.parametrize("fmt_name", ["NV21", "YV12"])
def test_quantized_model_inputs_rotated(O, angle, fmt_name):
    """Rotate -> CSC -> Resize -> Normalize(128, 1) into int8: the tile-staged kernel with the byte table."""
    W, H, d = 1280, 720, 224
    fmt = getattr(O, "FMT_" + fmt_name)
    ops = [(O.OP_ROTATE, angle), (O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, d, d), (O.OP_NORMALIZE, 0, 0, 0, 0, 128.0, 1.0)]
    fr = rand_frames(O, 2, W, H, fmt, 777 + angle)
    st, want = oracle_chain(O, ops, fr, W, H, fmt, O.DT_I8, (1, d, d, 3))
    assert st == 0
    st, got = gpu_chain(ops, fr, W, H, fmt, O.DT_I8, (1, d, d, 3), device_resident=True)
    assert st == 0
    assert_same_bits(got, want, "quantized rotated %s %d" % (fmt_name, angle))
