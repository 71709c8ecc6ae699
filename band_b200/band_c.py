"""ctypes binding of Band's C API as exported by lib/libband_c.so (include/band_c_api.h).

This is the harness-side view of the drop-in boundary: the same `Band*` calls a
C / JNI / dlopen user of the reference makes (band/c/c_api_buffer.h:55-121),
plus the additive device/batch entry points. No arithmetic happens in Python and
there is no fallback: a missing library or a missing CUDA device raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libband_c.so")

# c_api_type.h
kBandOk, kBandErr = 0, 1
(kBandNoType, kBandFloat32, kBandInt32, kBandUInt8, kBandInt64, kBandString, kBandBool, kBandInt16, kBandComplex64,
 kBandInt8, kBandFloat16, kBandFloat64) = range(12)
(kBandGrayScale, kBandRGB, kBandRGBA, kBandYV12, kBandYV21, kBandNV21, kBandNV12, kBandRaw) = range(8)
(BAND_CROP, BAND_RESIZE, BAND_ROTATE, BAND_FLIP, BAND_COLOR_SPACE_CONVERT, BAND_NORMALIZE,
 BAND_DATA_TYPE_CONVERT, BAND_AUTO_ORIENT) = range(8)
kBandHostMemory, kBandDeviceMemory = 0, 1

FORMATS = {"GRAY": kBandGrayScale, "RGB": kBandRGB, "RGBA": kBandRGBA, "YV12": kBandYV12, "YV21": kBandYV21,
           "NV21": kBandNV21, "NV12": kBandNV12}
NP_DTYPE = {kBandFloat32: np.float32, kBandInt32: np.int32, kBandUInt8: np.uint8, kBandInt16: np.int16,
            kBandInt8: np.int8, kBandFloat64: np.float64}
BAND_DTYPE = {np.dtype(v): k for k, v in NP_DTYPE.items()}

EXPORTED = [
    "BandTensorDelete", "BandTensorGetType", "BandTensorGetData", "BandTensorGetNumDims", "BandTensorGetDims",
    "BandTensorGetBytes", "BandTensorGetName", "BandTensorGetQuantizationType", "BandTensorGetQuantizationParams",
    "BandBufferCreate", "BandBufferDelete", "BandBufferSetFromRawData", "BandBufferSetFromYUVData",
    "BandImageProcessorBuilderCreate", "BandImageProcessorBuilderDelete", "BandImageProcessorBuilderBuild",
    "BandAddOperator", "BandImageProcessorProcess", "BandImageProcessorDelete",
    # additive
    "BandTensorCreate", "BandTensorGetMemorySpace", "BandBufferSetFromDeviceRawData",
    "BandBufferSetFromDeviceYUVData", "BandImageProcessorProcessBatch", "BandImageProcessorProcessBatchAsync",
    "BandImageProcessorSynchronize", "BandImageProcessorProcessBuffers", "BandImageProcessorGetStream", "BandImageProcessorIsFused",
    "BandCudaGetDeviceCount", "BandCudaSetDevice", "BandHostAlloc", "BandHostFree", "BandDeviceAlloc",
    "BandDeviceFree", "BandCopyToDevice", "BandCopyToHost", "BandCudaLaunchCount", "BandGetLastError",
    "BandTensorRingBufferCreate", "BandTensorRingBufferDelete", "BandTensorRingBufferAlloc",
    "BandTensorRingBufferIsHandleValid", "BandTensorRingBufferGetSlot", "BandTensorRingBufferPutTensor",
    "BandTensorRingBufferGetTensor", "BandBufferSetOrientation", "BandImageProcessorRecordBatch",
    "BandRecordedBatchReplay", "BandRecordedBatchDelete",
    "BandImageProcessorBuilderBuildSharded", "BandShardedImageProcessorDelete", "BandShardedImageProcessorGetNumShards",
    "BandShardedImageProcessorGetDevice", "BandShardedImageProcessorGetSlice", "BandImageProcessorProcessBatchSharded",
    "BandImageProcessorProcessBatchShardedAsync", "BandShardedImageProcessorSynchronize",
    "BandShardedImageProcessorGetStream", "BandShardedImageProcessorIsFused",
    "BandHostRegistrationCacheConfigure", "BandHostRegistrationCacheStats", "BandShardSlice",
]


class BandError(RuntimeError):
    pass


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise BandError("%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'`" % LIB_PATH)
    L = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    vp, sz, ci = C.c_void_p, C.c_size_t, C.c_int
    L.BandTensorDelete.argtypes = [vp]
    L.BandTensorGetType.argtypes = [vp]
    L.BandTensorGetData.argtypes = [vp]
    L.BandTensorGetData.restype = vp
    L.BandTensorGetNumDims.argtypes = [vp]
    L.BandTensorGetNumDims.restype = sz
    L.BandTensorGetDims.argtypes = [vp]
    L.BandTensorGetDims.restype = C.POINTER(ci)
    L.BandTensorGetBytes.argtypes = [vp]
    L.BandTensorGetBytes.restype = sz
    L.BandTensorGetName.argtypes = [vp]
    L.BandTensorGetName.restype = C.c_char_p
    L.BandTensorGetQuantizationType.argtypes = [vp]
    L.BandTensorGetQuantizationParams.argtypes = [vp]
    L.BandTensorGetQuantizationParams.restype = vp
    L.BandBufferCreate.restype = vp
    L.BandBufferDelete.argtypes = [vp]
    L.BandBufferSetFromRawData.argtypes = [vp, vp, sz, sz, ci]
    L.BandBufferSetFromYUVData.argtypes = [vp, vp, vp, vp, sz, sz, sz, sz, sz, ci]
    L.BandImageProcessorBuilderCreate.restype = vp
    L.BandImageProcessorBuilderDelete.argtypes = [vp]
    L.BandImageProcessorBuilderBuild.argtypes = [vp]
    L.BandImageProcessorBuilderBuild.restype = vp
    # BandAddOperator is variadic: argtypes are set per call site below
    L.BandImageProcessorProcess.argtypes = [vp, vp, vp]
    L.BandImageProcessorDelete.argtypes = [vp]
    L.BandTensorCreate.argtypes = [ci, sz, C.POINTER(ci), ci]
    L.BandTensorCreate.restype = vp
    L.BandTensorGetMemorySpace.argtypes = [vp]
    L.BandBufferSetFromDeviceRawData.argtypes = [vp, vp, sz, sz, ci]
    L.BandBufferSetFromDeviceYUVData.argtypes = [vp, vp, vp, vp, sz, sz, sz, sz, sz, ci]
    L.BandImageProcessorProcessBatch.argtypes = [vp, C.POINTER(vp), C.POINTER(vp), sz]
    L.BandImageProcessorProcessBatchAsync.argtypes = [vp, C.POINTER(vp), C.POINTER(vp), sz]
    L.BandImageProcessorSynchronize.argtypes = [vp]
    L.BandImageProcessorProcessBuffers.argtypes = [vp, C.POINTER(vp), C.POINTER(vp), sz, ci]
    L.BandImageProcessorGetStream.argtypes = [vp]
    L.BandImageProcessorGetStream.restype = vp
    L.BandImageProcessorIsFused.argtypes = [vp]
    L.BandCudaSetDevice.argtypes = [ci]
    L.BandHostAlloc.argtypes = [sz]
    L.BandHostAlloc.restype = vp
    L.BandHostFree.argtypes = [vp]
    L.BandDeviceAlloc.argtypes = [sz]
    L.BandDeviceAlloc.restype = vp
    L.BandDeviceFree.argtypes = [vp]
    L.BandCopyToDevice.argtypes = [vp, vp, sz]
    L.BandCopyToHost.argtypes = [vp, vp, sz]
    L.BandCudaLaunchCount.restype = C.c_longlong
    L.BandGetLastError.restype = C.c_char_p
    L.BandBufferSetOrientation.argtypes = [vp, ci]
    L.BandImageProcessorRecordBatch.argtypes = [vp, C.POINTER(vp), C.POINTER(vp), sz]
    L.BandImageProcessorRecordBatch.restype = vp
    L.BandRecordedBatchReplay.argtypes = [vp, ci]
    L.BandRecordedBatchDelete.argtypes = [vp]
    L.BandTensorRingBufferCreate.argtypes = [C.POINTER(vp), C.POINTER(ci), sz, ci, ci]
    L.BandTensorRingBufferCreate.restype = vp
    L.BandTensorRingBufferDelete.argtypes = [vp]
    L.BandTensorRingBufferAlloc.argtypes = [vp]
    L.BandTensorRingBufferIsHandleValid.argtypes = [vp, ci]
    L.BandTensorRingBufferGetSlot.argtypes = [vp, ci, ci]
    L.BandTensorRingBufferGetSlot.restype = vp
    L.BandTensorRingBufferPutTensor.argtypes = [vp, vp, ci, ci]
    L.BandTensorRingBufferGetTensor.argtypes = [vp, vp, ci, ci]
    L.BandImageProcessorBuilderBuildSharded.argtypes = [vp, C.POINTER(ci), sz]
    L.BandImageProcessorBuilderBuildSharded.restype = vp
    L.BandShardedImageProcessorDelete.argtypes = [vp]
    L.BandShardedImageProcessorGetNumShards.argtypes = [vp]
    L.BandShardedImageProcessorGetNumShards.restype = sz
    L.BandShardedImageProcessorGetDevice.argtypes = [vp, sz]
    L.BandShardedImageProcessorGetSlice.argtypes = [vp, sz, sz, C.POINTER(sz), C.POINTER(sz)]
    L.BandShardedImageProcessorGetSlice.restype = None
    L.BandImageProcessorProcessBatchSharded.argtypes = [vp, C.POINTER(vp), C.POINTER(vp), sz]
    L.BandImageProcessorProcessBatchShardedAsync.argtypes = [vp, C.POINTER(vp), C.POINTER(vp), sz]
    L.BandShardedImageProcessorSynchronize.argtypes = [vp]
    L.BandShardedImageProcessorGetStream.argtypes = [vp, sz]
    L.BandShardedImageProcessorGetStream.restype = vp
    L.BandShardedImageProcessorIsFused.argtypes = [vp, sz]
    L.BandShardSlice.argtypes = [sz, sz, sz, C.POINTER(sz), C.POINTER(sz)]
    L.BandShardSlice.restype = None
    L.BandHostRegistrationCacheConfigure.argtypes = [ci, sz]
    L.BandHostRegistrationCacheStats.argtypes = [C.POINTER(sz), C.POINTER(sz), C.POINTER(C.c_ulonglong),
                                                 C.POINTER(C.c_ulonglong)]
    L.BandHostRegistrationCacheStats.restype = None
    _lib = L
    return L


def last_error():
    return lib().BandGetLastError().decode(errors="replace")


def add_operator(builder, field, *args):
    """BandAddOperator(builder, field, count, ...): ints for image operators, doubles for Normalize."""
    L = lib()
    if field == BAND_NORMALIZE:
        cargs = [C.c_double(float(a)) for a in args]
    else:
        cargs = [C.c_int(int(a)) for a in args]
    return L.BandAddOperator(C.c_void_p(builder), C.c_int(field), C.c_int(len(args)), *cargs)


# ---------------------------------------------------------------------------------
# Thin RAII wrappers used by tests / bench / smoke. They only forward to the C API.
# ---------------------------------------------------------------------------------
class Tensor:
    """BandTensor created stand-alone (BandTensorCreate, additive)."""

    def __init__(self, np_dtype, dims, device=False):
        self.dims = tuple(int(d) for d in dims)
        self.np_dtype = np.dtype(np_dtype)
        arr = (C.c_int * len(self.dims))(*self.dims)
        self.device = bool(device)
        self.handle = lib().BandTensorCreate(BAND_DTYPE[self.np_dtype], len(self.dims), arr,
                                             kBandDeviceMemory if device else kBandHostMemory)
        if not self.handle:
            raise BandError("BandTensorCreate failed (no CUDA device or out of memory)")

    @property
    def data_ptr(self):
        return lib().BandTensorGetData(self.handle)

    @property
    def nbytes(self):
        return lib().BandTensorGetBytes(self.handle)

    def numpy(self):
        """Host copy of the tensor contents (downloads a device tensor)."""
        out = np.empty(self.dims, dtype=self.np_dtype)
        if self.device:
            if lib().BandCopyToHost(out.ctypes.data, self.data_ptr, self.nbytes) != kBandOk:
                raise BandError("BandCopyToHost failed")
        else:
            C.memmove(out.ctypes.data, self.data_ptr, self.nbytes)
        return out

    def close(self):
        if self.handle:
            lib().BandTensorDelete(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class SlotTensor(Tensor):
    """A ring slot lent by BandTensorRingBufferGetSlot: owned by the ring, never deleted here."""

    def __init__(self, handle, np_dtype, dims, device):
        self.handle = handle
        self.np_dtype = np.dtype(np_dtype)
        self.dims = tuple(int(d) for d in dims)
        self.device = bool(device)

    def close(self):
        self.handle = None


class TensorRingBuffer:
    """BandTensorRingBuffer (additive; band::TensorRingBuffer with host or HBM slots)."""

    def __init__(self, model_tensors, tensor_indices, size=128, device=False):
        n = len(model_tensors)
        self.model_tensors = list(model_tensors)
        self.indices = [int(i) for i in tensor_indices]
        self.device = bool(device)
        ta = (C.c_void_p * n)(*[t.handle for t in model_tensors])
        ia = (C.c_int * n)(*self.indices)
        self.handle = lib().BandTensorRingBufferCreate(ta, ia, n, int(size),
                                                       kBandDeviceMemory if device else kBandHostMemory)
        if not self.handle:
            raise BandError("BandTensorRingBufferCreate failed: " + last_error())

    def alloc(self):
        return lib().BandTensorRingBufferAlloc(self.handle)

    def is_handle_valid(self, handle):
        return bool(lib().BandTensorRingBufferIsHandleValid(self.handle, handle))

    def slot(self, tensor_index, handle):
        h = lib().BandTensorRingBufferGetSlot(self.handle, tensor_index, handle)
        if not h:
            return None
        like = self.model_tensors[self.indices.index(tensor_index)]
        return SlotTensor(h, like.np_dtype, like.dims, self.device)

    def put(self, src, tensor_index, handle):
        return lib().BandTensorRingBufferPutTensor(self.handle, src.handle, tensor_index, handle)

    def get(self, dst, tensor_index, handle):
        return lib().BandTensorRingBufferGetTensor(self.handle, dst.handle, tensor_index, handle)

    def close(self):
        if self.handle:
            lib().BandTensorRingBufferDelete(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Buffer:
    """BandBuffer: a non-owning view; `keep` pins the Python object that owns the bytes."""

    def __init__(self):
        self.handle = lib().BandBufferCreate()
        self.keep = None

    def set_raw(self, ptr, width, height, fmt, device=False, keep=None):
        f = lib().BandBufferSetFromDeviceRawData if device else lib().BandBufferSetFromRawData
        if f(self.handle, ptr, width, height, fmt) != kBandOk:
            raise BandError("BandBufferSetFromRawData failed")
        self.keep = keep
        return self

    def set_yuv(self, y, u, v, width, height, row_stride_y, row_stride_uv, pixel_stride_uv, fmt, device=False,
                keep=None):
        f = lib().BandBufferSetFromDeviceYUVData if device else lib().BandBufferSetFromYUVData
        if f(self.handle, y, u, v, width, height, row_stride_y, row_stride_uv, pixel_stride_uv, fmt) != kBandOk:
            raise BandError("BandBufferSetFromYUVData failed")
        self.keep = keep
        return self

    def set_orientation(self, exif):
        if lib().BandBufferSetOrientation(self.handle, int(exif)) != kBandOk:
            raise BandError("BandBufferSetOrientation failed")
        return self

    @staticmethod
    def from_numpy(frame, width, height, fmt):
        frame = np.ascontiguousarray(frame)
        return Buffer().set_raw(frame.ctypes.data, width, height, fmt, device=False, keep=frame)

    def close(self):
        if self.handle:
            lib().BandBufferDelete(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class ImageProcessor:
    """BandImageProcessorBuilder + Build. ops: [(field, *args)], e.g.
    [(BAND_COLOR_SPACE_CONVERT, kBandRGB), (BAND_RESIZE, 224, 224), (BAND_NORMALIZE, 127.5, 127.5)]."""

    def __init__(self, ops):
        L = lib()
        builder = L.BandImageProcessorBuilderCreate()
        try:
            for op in ops:
                if add_operator(builder, op[0], *op[1:]) != kBandOk:
                    raise BandError("BandAddOperator rejected %r" % (op,))
            self.handle = L.BandImageProcessorBuilderBuild(builder)
        finally:
            L.BandImageProcessorBuilderDelete(builder)
        if not self.handle:
            raise BandError("BandImageProcessorBuilderBuild failed: %s" % last_error())

    def process(self, buffer, tensor):
        """BandImageProcessorProcess; returns the BandStatus."""
        return lib().BandImageProcessorProcess(self.handle, buffer.handle, tensor.handle)

    def process_batch(self, buffers, tensors):
        n = len(buffers)
        ba = (C.c_void_p * n)(*[b.handle for b in buffers])
        ta = (C.c_void_p * n)(*[t.handle for t in tensors])
        return lib().BandImageProcessorProcessBatch(self.handle, ba, ta, n)

    def record_batch(self, buffers, tensors):
        """BandImageProcessorRecordBatch -> opaque handle (replay with replay(), free with release())."""
        n = len(buffers)
        ba = (C.c_void_p * n)(*[b.handle for b in buffers])
        ta = (C.c_void_p * n)(*[t.handle for t in tensors])
        h = lib().BandImageProcessorRecordBatch(self.handle, ba, ta, n)
        if not h:
            raise BandError("BandImageProcessorRecordBatch failed: " + last_error())
        return h

    @staticmethod
    def replay(recorded, synchronize=True):
        return lib().BandRecordedBatchReplay(recorded, 1 if synchronize else 0)

    @staticmethod
    def release(recorded):
        lib().BandRecordedBatchDelete(recorded)

    def process_batch_handles(self, buffer_array, tensor_array, n):
        """Same call with pre-built ctypes arrays (keeps Python overhead out of timed loops)."""
        return lib().BandImageProcessorProcessBatch(self.handle, buffer_array, tensor_array, n)

    def process_batch_async(self, buffer_array, tensor_array, n):
        return lib().BandImageProcessorProcessBatchAsync(self.handle, buffer_array, tensor_array, n)

    def synchronize(self):
        return lib().BandImageProcessorSynchronize(self.handle)

    @property
    def stream(self):
        return lib().BandImageProcessorGetStream(self.handle)

    @property
    def is_fused(self):
        return bool(lib().BandImageProcessorIsFused(self.handle))

    def close(self):
        if self.handle:
            lib().BandImageProcessorDelete(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class ShardedImageProcessor:
    """BandImageProcessorBuilderBuildSharded: one processor (plans, streams, host thread) per device; a batch is
    cut into contiguous slices, one per shard, by ONE call. devices=None: every visible GPU."""

    def __init__(self, ops, devices=None):
        L = lib()
        builder = L.BandImageProcessorBuilderCreate()
        try:
            for op in ops:
                if add_operator(builder, op[0], *op[1:]) != kBandOk:
                    raise BandError("BandAddOperator rejected %r" % (op,))
            if devices is None:
                self.handle = L.BandImageProcessorBuilderBuildSharded(builder, None, 0)
            else:
                arr = (C.c_int * len(devices))(*[int(d) for d in devices])
                self.handle = L.BandImageProcessorBuilderBuildSharded(builder, arr, len(devices))
        finally:
            L.BandImageProcessorBuilderDelete(builder)
        if not self.handle:
            raise BandError("BandImageProcessorBuilderBuildSharded failed: %s" % last_error())

    @property
    def num_shards(self):
        return int(lib().BandShardedImageProcessorGetNumShards(self.handle))

    def device_of(self, shard):
        return int(lib().BandShardedImageProcessorGetDevice(self.handle, shard))

    def slice_of(self, count, shard):
        b, e = C.c_size_t(), C.c_size_t()
        lib().BandShardedImageProcessorGetSlice(self.handle, count, shard, C.byref(b), C.byref(e))
        return int(b.value), int(e.value)

    def process_batch(self, buffers, tensors):
        n = len(buffers)
        ba = (C.c_void_p * n)(*[b.handle for b in buffers])
        ta = (C.c_void_p * n)(*[t.handle for t in tensors])
        return lib().BandImageProcessorProcessBatchSharded(self.handle, ba, ta, n)

    def process_batch_handles(self, buffer_array, tensor_array, n):
        return lib().BandImageProcessorProcessBatchSharded(self.handle, buffer_array, tensor_array, n)

    def process_batch_async(self, buffer_array, tensor_array, n):
        return lib().BandImageProcessorProcessBatchShardedAsync(self.handle, buffer_array, tensor_array, n)

    def synchronize(self):
        return lib().BandShardedImageProcessorSynchronize(self.handle)

    def stream(self, shard):
        return lib().BandShardedImageProcessorGetStream(self.handle, shard)

    def is_fused(self, shard):
        return bool(lib().BandShardedImageProcessorIsFused(self.handle, shard))

    def close(self):
        if self.handle:
            lib().BandShardedImageProcessorDelete(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def shard_slice(count, shards, shard):
    """BandShardSlice: the contiguous-slice rule of the sharded processor."""
    b, e = C.c_size_t(), C.c_size_t()
    lib().BandShardSlice(count, shards, shard, C.byref(b), C.byref(e))
    return int(b.value), int(e.value)


def host_registration_cache(enable, max_bytes=0):
    """BandHostRegistrationCacheConfigure: processors built afterwards register pageable frames on first sight."""
    return lib().BandHostRegistrationCacheConfigure(1 if enable else 0, int(max_bytes))


def host_registration_stats():
    b, e = C.c_size_t(), C.c_size_t()
    h, r = C.c_ulonglong(), C.c_ulonglong()
    lib().BandHostRegistrationCacheStats(C.byref(b), C.byref(e), C.byref(h), C.byref(r))
    return {"registered_bytes": int(b.value), "entries": int(e.value), "hits": int(h.value),
            "registrations": int(r.value)}


def frame_bytes(width, height, fmt):
    """Bytes of one contiguous frame (Buffer::GetBufferByteSize, buffer.cc:233-266)."""
    if fmt in (kBandNV12, kBandNV21, kBandYV12, kBandYV21):
        return width * height + 2 * ((width + 1) // 2) * ((height + 1) // 2)
    return width * height * {kBandGrayScale: 1, kBandRGB: 3, kBandRGBA: 4}[fmt]


def run_chain(ops, frames, width, height, fmt, out_dtype, out_dims, device_resident=False):
    """Convenience for tests/smoke: n contiguous host frames (n, frame_bytes) through one processor.

    device_resident=True uploads the frames first and writes device tensors (the
    HBM-resident path); otherwise host buffers go straight into the C API.
    Returns (status, ndarray (n, *out_dims))."""
    L = lib()
    frames = np.ascontiguousarray(frames, dtype=np.uint8)
    n = frames.shape[0]
    proc = ImageProcessor(ops)
    tensors = [Tensor(out_dtype, out_dims, device=device_resident) for _ in range(n)]
    bufs, dev = [], None
    try:
        if device_resident:
            dev = L.BandDeviceAlloc(frames.nbytes)
            if not dev or L.BandCopyToDevice(dev, frames.ctypes.data, frames.nbytes) != kBandOk:
                raise BandError("device upload failed")
            for i in range(n):
                bufs.append(Buffer().set_raw(dev + i * frames.shape[1], width, height, fmt, device=True))
        else:
            for i in range(n):
                bufs.append(Buffer().set_raw(frames[i].ctypes.data, width, height, fmt, keep=frames))
        st = proc.process_batch(bufs, tensors) if n != 1 else proc.process(bufs[0], tensors[0])
        out = np.stack([t.numpy() for t in tensors]) if st == kBandOk else None
        return st, out
    finally:
        for b in bufs:
            b.close()
        for t in tensors:
            t.close()
        proc.close()
        if dev:
            L.BandDeviceFree(dev)
