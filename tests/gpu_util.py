"""Shared helpers of the GPU parity tests: run a chain through the product's C API
(libband_c.so) and through the CPU oracle on the same seeded input."""
import ctypes as C
import os

import numpy as np

from conftest import ROOT

# oracle op kind -> C-API field is the identity for 0..6 (c_api_type.h:178-186); 7 = AutoConvert = empty builder


def to_capi_ops(ops):
    """Oracle-style op tuples -> band_c.ImageProcessor op tuples."""
    from band_b200 import band_c as B
    out = []
    for op in ops:
        k = op[0]
        if k == 7:
            continue  # AutoConvert: the builder's default when no operator is added
        if k == B.BAND_NORMALIZE:
            out.append((k, op[5], op[6]))
        elif k == B.BAND_CROP:
            out.append((k,) + tuple(op[1:5]))
        elif k in (B.BAND_RESIZE, B.BAND_FLIP):
            out.append((k,) + tuple(op[1:3]))
        elif k in (B.BAND_ROTATE, B.BAND_COLOR_SPACE_CONVERT):
            out.append((k, op[1]))
        else:
            out.append((k,))
    return out


NP_OF_ORC = {1: np.float32, 2: np.int32, 3: np.uint8, 7: np.int16, 9: np.int8, 11: np.float64}


def gpu_chain(ops, frames, W, H, fmt, out_dtype, out_dims, device_resident=False):
    """(status, ndarray (n,*out_dims)) through BandImageProcessorProcess[Batch]."""
    from band_b200 import band_c as B
    frames = np.ascontiguousarray(frames, dtype=np.uint8)
    if frames.ndim == 1:
        frames = frames[None]
    return B.run_chain(to_capi_ops(ops), frames, W, H, fmt, NP_OF_ORC[out_dtype], out_dims,
                       device_resident=device_resident)


def oracle_chain(O, ops, frames, W, H, fmt, out_dtype, out_dims):
    frames = np.ascontiguousarray(frames, dtype=np.uint8)
    if frames.ndim == 1:
        frames = frames[None]
    outs = []
    for f in frames:
        st, out = O.run_chain(ops, f, W, H, fmt, out_dtype, out_dims)
        if st != 0:
            return st, None
        outs.append(out)
    return 0, np.stack(outs)


def assert_same_bits(got, want, what=""):
    assert got is not None and want is not None, what
    assert got.shape == want.shape and got.dtype == want.dtype, (what, got.shape, want.shape, got.dtype, want.dtype)
    g = np.ascontiguousarray(got).view(np.uint8)
    w = np.ascontiguousarray(want).view(np.uint8)
    bad = int((g != w).sum())
    assert bad == 0, "%s: %d of %d bytes differ" % (what, bad, g.size)


def rand_frames(O, n, W, H, fmt, seed):
    rng = np.random.default_rng(seed)
    return rng.integers(0, 256, size=(n, O.frame_bytes(W, H, fmt)), dtype=np.uint8)


# ---- the reference tests' C++ call sequences, via tests/cc/band_testing.cc -------------------------
class OpDesc(C.Structure):
    _fields_ = [("kind", C.c_int), ("i", C.c_int * 4), ("f", C.c_float * 2), ("inplace", C.c_int)]


class OutInfo(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("format", C.c_int), ("dtype", C.c_int),
                ("bytes", C.c_size_t)]


_T = None


def testing_lib():
    global _T
    if _T is None:
        from band_b200 import band_c as B
        B.lib()
        T = C.CDLL(os.path.join(ROOT, "band_b200", "lib", "libband_testing.so"))
        T.bandt_run_op.argtypes = [C.POINTER(OpDesc), C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                   C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(OutInfo)]
        T.bandt_run_chain.argtypes = [C.POINTER(OpDesc), C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                      C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(OutInfo),
                                      C.POINTER(C.c_int)]
        _T = T
    return _T


def make_desc(op):
    d = OpDesc()
    d.kind = op[0]
    for j, v in enumerate(list(op[1:5])):
        d.i[j] = int(v)
    for j, v in enumerate(list(op[5:7])):
        d.f[j] = float(v)
    d.inplace = int(op[7]) if len(op) > 7 else 0
    return d


def gpu_op(op, frame, W, H, fmt, in_dtype=3, assign=None, cap=None):
    """IBufferOperator::Process stand-alone on the GPU backing. assign = (format, dtype, w, h) or None.
    Returns (absl status code, bytes ndarray, (w, h, format, dtype))."""
    T = testing_lib()
    frame = np.ascontiguousarray(frame)
    cap = cap or max(64, frame.nbytes * 8)
    out = np.zeros(cap, dtype=np.uint8)
    info = OutInfo()
    d = make_desc(op)
    a = assign or (-1, 3, 0, 0)
    st = T.bandt_run_op(C.byref(d), frame.ctypes.data, W, H, fmt, in_dtype, a[0], a[1], a[2], a[3],
                        out.ctypes.data, cap, C.byref(info))
    if st != 0:
        return st, None, None
    assert info.bytes <= cap
    return 0, out[:info.bytes].copy(), (info.width, info.height, info.format, info.dtype)


def gpu_image_chain(ops, frame, W, H, fmt, out_format, out_dtype, out_w, out_h):
    """ImageProcessorBuilder -> Process(input, image output buffer). Returns (status, bytes, info)."""
    T = testing_lib()
    frame = np.ascontiguousarray(frame)
    arr = (OpDesc * max(1, len(ops)))()
    for k, op in enumerate(ops):
        arr[k] = make_desc(op)
    cap = max(64, out_w * out_h * 16)
    out = np.zeros(cap, dtype=np.uint8)
    info = OutInfo()
    bst = C.c_int(0)
    st = T.bandt_run_chain(arr, len(ops), frame.ctypes.data, W, H, fmt, out_format, out_dtype, out_w, out_h,
                           out.ctypes.data, cap, C.byref(info), C.byref(bst))
    if st != 0:
        return st, None, None
    return 0, out[:info.bytes].copy(), (info.width, info.height, info.format, info.dtype)
