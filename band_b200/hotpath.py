"""The hot path as a user of the reference drives it, through Band's C API (lib/libband_c.so):

    BandImageProcessorBuilder + BandAddOperator(COLOR_SPACE_CONVERT / RESIZE / NORMALIZE ...)
    -> BandImageProcessorBuilderBuild -> BandImageProcessorProcess[Batch](buffers, tensors)

(/root/reference/band/c/c_api_buffer.h:55-102; the C++ call sequence is
band/test/backend/tfl_minimal_test.cc:504-515). Nothing here computes pixels and nothing
falls back to a CPU implementation: every call lands in the fused CUDA plan, or raises.

Two helpers are used by smoke(), the tests and bench.py:
  * run_host_batch()   host frames in, host result out (uploads and downloads inside the call);
  * ResidentBatch      frames and tensors resident in HBM, one ProcessBatch per step.
"""
import ctypes as C

import numpy as np

from . import band_c as B

FORMAT = B.FORMATS  # "NV12" -> kBandNV12 ...


def chain_ops(dst_w, dst_h, out_f32, mean=127.5, std=127.5, to_format="RGB", crop=None, rotate=None, flip=None):
    """Operator list in the order the reference's users add them (Rotate, Crop, ColorSpaceConvert, Resize, Normalize)."""
    ops = []
    if rotate:
        ops.append((B.BAND_ROTATE, int(rotate)))
    if crop:
        ops.append((B.BAND_CROP,) + tuple(int(c) for c in crop))
    if flip:
        ops.append((B.BAND_FLIP, int(flip[0]), int(flip[1])))
    if to_format is not None:
        ops.append((B.BAND_COLOR_SPACE_CONVERT, FORMAT[to_format]))
    if dst_w and dst_h:
        ops.append((B.BAND_RESIZE, int(dst_w), int(dst_h)))
    if out_f32:
        ops.append((B.BAND_NORMALIZE, float(mean), float(std)))
    return ops


def set_device(device):
    if B.lib().BandCudaSetDevice(int(device)) != B.kBandOk:
        raise B.BandError("BandCudaSetDevice(%d) failed: %s" % (device, B.last_error()))


def launch_count():
    """Kernel launches issued by libbandcu.so in this process."""
    return int(B.lib().BandCudaLaunchCount())


def run_host_batch(frames, width, height, fmt, dst_w, dst_h, out_f32=True, mean=127.5, std=127.5, device=0,
                   channels=3, **chain):
    """n host frames (n, frame_bytes) uint8 -> ndarray (n, 1, dst_h, dst_w, channels), via host BandBuffers and
    host BandTensors (the reference-facing call; H2D/D2H happen inside BandImageProcessorProcessBatch)."""
    set_device(device)
    ops = chain_ops(dst_w, dst_h, out_f32, mean, std, **chain)
    dt = np.float32 if out_f32 else np.uint8
    st, out = B.run_chain(ops, frames, width, height, FORMAT[fmt], dt, (1, dst_h, dst_w, channels))
    if st != B.kBandOk:
        raise B.BandError("BandImageProcessorProcess failed (%d): %s" % (st, B.last_error()))
    return out


class ResidentBatch:
    """n frames of one layout resident in HBM plus n device tensors and one processor.

    step() is exactly one BandImageProcessorProcessBatch over the resident buffers. Used for the
    device-resident throughput figure (bench.py `value`) and by the GPU parity tests.
    """

    def __init__(self, ops, n, width, height, fmt, out_dtype, out_dims, device=0):
        set_device(device)
        L = B.lib()
        self.n, self.width, self.height = int(n), int(width), int(height)
        self.fmt = FORMAT[fmt] if isinstance(fmt, str) else int(fmt)
        self.frame_bytes = B.frame_bytes(self.width, self.height, self.fmt)
        self.out_dims = tuple(out_dims)
        self.out_dtype = np.dtype(out_dtype)
        self.out_bytes = int(np.prod(self.out_dims)) * self.out_dtype.itemsize
        self.proc = B.ImageProcessor(ops)
        # frames 256-byte aligned so the TMA staging path applies to every frame
        self.frame_pitch = (self.frame_bytes + 255) // 256 * 256
        self.dev_frames = L.BandDeviceAlloc(self.frame_pitch * self.n)
        if not self.dev_frames:
            raise B.BandError("BandDeviceAlloc failed: %s" % B.last_error())
        self.tensors = [B.Tensor(self.out_dtype, self.out_dims, device=True) for _ in range(self.n)]
        self.buffers = [B.Buffer().set_raw(self.dev_frames + i * self.frame_pitch, self.width, self.height, self.fmt,
                                           device=True) for i in range(self.n)]
        self._ba = (C.c_void_p * self.n)(*[b.handle for b in self.buffers])
        self._ta = (C.c_void_p * self.n)(*[t.handle for t in self.tensors])

    def upload(self, frames):
        """frames: (n, frame_bytes) uint8 host array."""
        frames = np.ascontiguousarray(frames, dtype=np.uint8)
        assert frames.shape == (self.n, self.frame_bytes), frames.shape
        L = B.lib()
        if self.frame_pitch == self.frame_bytes:
            ok = L.BandCopyToDevice(self.dev_frames, frames.ctypes.data, frames.nbytes) == B.kBandOk
        else:
            ok = all(L.BandCopyToDevice(self.dev_frames + i * self.frame_pitch, frames[i].ctypes.data,
                                        self.frame_bytes) == B.kBandOk for i in range(self.n))
        if not ok:
            raise B.BandError("BandCopyToDevice failed: %s" % B.last_error())

    def step(self):
        st = self.proc.process_batch_handles(self._ba, self._ta, self.n)
        if st != B.kBandOk:
            raise B.BandError("BandImageProcessorProcessBatch failed (%d): %s" % (st, B.last_error()))

    def download(self, indices=None):
        idx = range(self.n) if indices is None else indices
        return np.stack([self.tensors[i].numpy() for i in idx])

    @property
    def is_fused(self):
        return self.proc.is_fused

    def close(self):
        for b in getattr(self, "buffers", []):
            b.close()
        for t in getattr(self, "tensors", []):
            t.close()
        if getattr(self, "proc", None):
            self.proc.close()
        if getattr(self, "dev_frames", None):
            B.lib().BandDeviceFree(self.dev_frames)
            self.dev_frames = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class ShardedResidentBatch:
    """n frames of one layout spread over the shards of a BandShardedImageProcessor: every shard's contiguous slice
    (BandShardedImageProcessorGetSlice) is allocated on that shard's device together with its tensors, so ONE
    BandImageProcessorProcessBatchSharded[Async] call drives all devices with no peer traffic."""

    def __init__(self, ops, n, width, height, fmt, out_dtype, out_dims, devices, unique_frames, resident=True):
        L = B.lib()
        self.n, self.width, self.height = int(n), int(width), int(height)
        self.fmt = FORMAT[fmt] if isinstance(fmt, str) else int(fmt)
        self.frame_bytes = B.frame_bytes(self.width, self.height, self.fmt)
        self.out_dims, self.out_dtype = tuple(out_dims), np.dtype(out_dtype)
        self.out_bytes = int(np.prod(self.out_dims)) * self.out_dtype.itemsize
        self.resident = bool(resident)
        self.proc = B.ShardedImageProcessor(ops, devices)
        self.frame_pitch = (self.frame_bytes + 255) // 256 * 256
        self.buffers, self.tensors, self._allocs = [None] * self.n, [None] * self.n, []
        uniq = np.ascontiguousarray(unique_frames, dtype=np.uint8)
        for k in range(self.proc.num_shards):
            b, e = self.proc.slice_of(self.n, k)
            if e <= b:
                continue
            set_device(self.proc.device_of(k))
            if self.resident:
                base = L.BandDeviceAlloc(self.frame_pitch * (e - b))
                if not base:
                    raise B.BandError("BandDeviceAlloc failed: %s" % B.last_error())
                self._allocs.append(("dev", base))
            else:  # pinned host frames: read in place over PCIe by the shard's kernel
                base = L.BandHostAlloc(self.frame_pitch * (e - b))
                if not base:
                    raise B.BandError("BandHostAlloc failed")
                self._allocs.append(("host", base))
            for i in range(b, e):
                p = base + (i - b) * self.frame_pitch
                src = uniq[i % uniq.shape[0]]
                if self.resident:
                    if L.BandCopyToDevice(p, src.ctypes.data, self.frame_bytes) != B.kBandOk:
                        raise B.BandError("upload failed: " + B.last_error())
                else:
                    C.memmove(p, src.ctypes.data, self.frame_bytes)
                self.buffers[i] = B.Buffer().set_raw(p, self.width, self.height, self.fmt, device=self.resident)
                self.tensors[i] = B.Tensor(self.out_dtype, self.out_dims, device=self.resident)
        self._ba = (C.c_void_p * self.n)(*[b.handle for b in self.buffers])
        self._ta = (C.c_void_p * self.n)(*[t.handle for t in self.tensors])

    def step(self):
        if self.proc.process_batch_handles(self._ba, self._ta, self.n) != B.kBandOk:
            raise B.BandError("BandImageProcessorProcessBatchSharded failed: %s" % B.last_error())

    def step_async(self):
        if self.proc.process_batch_async(self._ba, self._ta, self.n) != B.kBandOk:
            raise B.BandError("BandImageProcessorProcessBatchShardedAsync failed: %s" % B.last_error())

    def synchronize(self):
        if self.proc.synchronize() != B.kBandOk:
            raise B.BandError("BandShardedImageProcessorSynchronize failed: %s" % B.last_error())

    def download(self, indices):
        return np.stack([self.tensors[i].numpy() for i in indices])

    def close(self):
        for b in getattr(self, "buffers", []):
            if b:
                b.close()
        for t in getattr(self, "tensors", []):
            if t:
                t.close()
        if getattr(self, "proc", None):
            self.proc.close()
            self.proc = None
        for kind, p in getattr(self, "_allocs", []):
            (B.lib().BandDeviceFree if kind == "dev" else B.lib().BandHostFree)(p)
        self._allocs = []
