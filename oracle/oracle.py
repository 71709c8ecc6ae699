"""ctypes front-end of the CPU oracle (TEST INFRASTRUCTURE ONLY).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this module. The product path (band_b200) never does.

The oracle restates the reference's band/buffer chain semantics
(/root/reference/band/buffer/*.cc) on top of a restatement of libyuv's C paths;
see oracle/yuvref.h for the provenance and pinning statement.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "liborc.so")

# band/common.h:115-161
FMT_GRAY, FMT_RGB, FMT_RGBA, FMT_YV12, FMT_YV21, FMT_NV21, FMT_NV12, FMT_RAW = range(8)
DT_F32, DT_I32, DT_U8, DT_I16, DT_I8, DT_F64 = 1, 2, 3, 7, 9, 11
OK, UNKNOWN, INVALID_ARGUMENT, INTERNAL = 0, 2, 3, 13
# c_api_type.h:178-186 (+ AutoConvert, which has no C-API field)
OP_CROP, OP_RESIZE, OP_ROTATE, OP_FLIP, OP_CSC, OP_NORMALIZE, OP_DTC, OP_AUTO = range(8)

_NP_DT = {DT_U8: np.uint8, DT_I8: np.int8, DT_I16: np.int16, DT_I32: np.int32, DT_F32: np.float32,
          DT_F64: np.float64}


def build(force=False):
    """Compile oracle/_build/liborc.so with the committed Makefile."""
    if force:
        subprocess.run(["make", "-C", _HERE, "clean"], check=True, capture_output=True)
    arch = "x86-64-v3"
    try:
        with open("/proc/cpuinfo") as f:
            if " avx2" not in f.read():
                arch = "x86-64"
    except OSError:
        pass
    subprocess.run(["make", "-C", _HERE, "ORC_ARCH=" + arch], check=True, capture_output=True)


class OpDesc(C.Structure):
    _fields_ = [("kind", C.c_int), ("i", C.c_int * 4), ("f", C.c_float * 2), ("inplace", C.c_int)]


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        build()
    else:
        try:
            with open("/proc/cpuinfo") as f:
                if " avx2" not in f.read():
                    build(force=True)
        except OSError:
            pass
    L = C.CDLL(_LIB_PATH)
    vp, sz, ci = C.c_void_p, C.c_size_t, C.c_int
    L.orc_buffer_from_raw.restype = vp
    L.orc_buffer_from_raw.argtypes = [vp, sz, sz, ci, ci]
    L.orc_buffer_from_yuv_planes.restype = vp
    L.orc_buffer_from_yuv_planes.argtypes = [vp, vp, vp, sz, sz, sz, sz, sz, ci]
    L.orc_buffer_from_tensor.restype = vp
    L.orc_buffer_from_tensor.argtypes = [vp, ci, C.POINTER(ci), ci]
    L.orc_buffer_empty.restype = vp
    L.orc_buffer_empty.argtypes = [sz, sz, ci, ci]
    L.orc_buffer_free.argtypes = [vp]
    L.orc_buffer_byte_size.restype = sz
    L.orc_buffer_byte_size.argtypes = [sz, sz, ci]
    L.orc_op_create.restype = vp
    L.orc_op_create.argtypes = [ci, ci, ci, ci, ci, C.c_float, C.c_float, ci]
    L.orc_op_free.argtypes = [vp]
    L.orc_op_process.argtypes = [vp, vp]
    L.orc_op_set_output.argtypes = [vp, vp]
    L.orc_op_output.restype = vp
    L.orc_op_output.argtypes = [vp]
    L.orc_processor_build.restype = vp
    L.orc_processor_build.argtypes = [C.POINTER(OpDesc), ci, C.POINTER(ci)]
    L.orc_processor_process.argtypes = [vp, vp, vp]
    L.orc_processor_free.argtypes = [vp]
    L.orc_run_batch.argtypes = [C.POINTER(OpDesc), ci, vp, sz, sz, sz, ci, vp, sz, ci, C.POINTER(ci), ci, ci]
    L.orc_YuvPixel.argtypes = [C.c_uint8, C.c_uint8, C.c_uint8, vp, vp, vp]
    _lib = L
    return L


class _Plane(C.Structure):
    _fields_ = [("data", C.c_void_p), ("row_stride_bytes", C.c_size_t), ("pixel_stride_bytes", C.c_size_t)]


class _Buffer(C.Structure):
    _fields_ = [("format", C.c_int), ("dtype", C.c_int), ("orientation", C.c_int), ("width", C.c_size_t),
                ("height", C.c_size_t), ("raw_ndims", C.c_size_t), ("raw_dims", C.c_size_t * 8),
                ("nplanes", C.c_int), ("planes", _Plane * 3), ("owns_data", C.c_int)]


def buffer_view(ptr):
    return C.cast(ptr, C.POINTER(_Buffer)).contents


def make_descs(ops):
    """ops: list of tuples (kind, i0, i1, i2, i3, f0, f1[, inplace]) or shorter; missing fields are 0."""
    arr = (OpDesc * max(1, len(ops)))()
    for k, op in enumerate(ops):
        op = tuple(op)
        arr[k].kind = op[0]
        ints = list(op[1:5]) + [0] * (4 - len(op[1:5]))
        for j in range(4):
            arr[k].i[j] = int(ints[j])
        fl = list(op[5:7]) + [0.0] * (2 - len(op[5:7]))
        arr[k].f[0], arr[k].f[1] = float(fl[0]), float(fl[1])
        arr[k].inplace = int(op[7]) if len(op) > 7 else 0
    return arr


def frame_bytes(width, height, fmt):
    return lib().orc_buffer_byte_size(width, height, fmt)


def run_chain(ops, frame, width, height, fmt, out_dtype, out_dims, in_dtype=DT_U8):
    """BufferProcessor::Process on one contiguous frame -> (status, ndarray shaped out_dims)."""
    L = lib()
    frame = np.ascontiguousarray(frame)
    out = np.zeros(out_dims, dtype=_NP_DT[out_dtype])
    descs = make_descs(ops)
    st = C.c_int(0)
    proc = L.orc_processor_build(descs, len(ops), C.byref(st))
    if not proc:
        return st.value, out
    dims = (C.c_int * len(out_dims))(*out_dims)
    ib = L.orc_buffer_from_raw(frame.ctypes.data, width, height, fmt, in_dtype)
    ob = L.orc_buffer_from_tensor(out.ctypes.data, out_dtype, dims, len(out_dims))
    try:
        status = L.orc_processor_process(proc, ib, ob) if ib and ob else INTERNAL
    finally:
        L.orc_buffer_free(ib)
        L.orc_buffer_free(ob)
        L.orc_processor_free(proc)
    return status, out


def run_op(op, frame, width, height, fmt, in_dtype=DT_U8):
    """IBufferOperator::Process standalone with an auto-created output.

    Returns (status, bytes ndarray of the whole output buffer, (w, h, format))."""
    L = lib()
    frame = np.ascontiguousarray(frame)
    d = make_descs([op])[0]
    o = L.orc_op_create(d.kind, d.i[0], d.i[1], d.i[2], d.i[3], d.f[0], d.f[1], d.inplace)
    ib = L.orc_buffer_from_raw(frame.ctypes.data, width, height, fmt, in_dtype)
    try:
        st = L.orc_op_process(o, ib)
        if st != OK:
            return st, None, None
        ob = buffer_view(L.orc_op_output(o))
        n = L.orc_buffer_byte_size(ob.width, ob.height, ob.format)
        raw = np.ctypeslib.as_array(C.cast(ob.planes[0].data, C.POINTER(C.c_uint8)), shape=(n,)).copy()
        return st, raw, (ob.width, ob.height, ob.format)
    finally:
        L.orc_buffer_free(ib)
        L.orc_op_free(o)


def run_batch(ops, frames, width, height, fmt, out_dtype, out_dims):
    """n frames back to back -> ndarray (n, *out_dims); the timed CPU-baseline entry."""
    L = lib()
    frames = np.ascontiguousarray(frames)
    fb = frame_bytes(width, height, fmt)
    n = frames.size // fb
    out = np.zeros((n,) + tuple(out_dims), dtype=_NP_DT[out_dtype])
    descs = make_descs(ops)
    dims = (C.c_int * len(out_dims))(*out_dims)
    st = L.orc_run_batch(descs, len(ops), frames.ctypes.data, fb, width, height, fmt, out.ctypes.data,
                         out[0].nbytes, out_dtype, dims, len(out_dims), n)
    return st, out
