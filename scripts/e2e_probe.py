#!/usr/bin/env python
"""Where does the host->host e2e time go? Times the north-star chain through the C API with every
combination of host/device inputs and outputs (pinned host memory)."""
import ctypes as C, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from band_b200 import band_c as B, hotpath

W, H, DW, DH = 1920, 1080, 224, 224
FB = W * H * 3 // 2
E = int(os.environ.get("E", "256"))
L = B.lib()
hotpath.set_device(0)
ops = hotpath.chain_ops(DW, DH, True)
host = L.BandHostAlloc(E * FB)
hv = np.ctypeslib.as_array(C.cast(host, C.POINTER(C.c_uint8)), shape=(E, FB))
hv[:] = np.random.default_rng(1).integers(0, 256, size=(1, FB), dtype=np.uint8)
dev = L.BandDeviceAlloc(E * FB)
L.BandCopyToDevice(dev, host, E * FB)
for in_dev in (False, True):
    for out_dev in (False, True):
        proc = B.ImageProcessor(ops)
        tens = [B.Tensor(np.float32, (1, DH, DW, 3), device=out_dev) for _ in range(E)]
        base = dev if in_dev else host
        bufs = [B.Buffer().set_raw(base + i * FB, W, H, B.kBandNV12, device=in_dev) for i in range(E)]
        ba = (C.c_void_p * E)(*[b.handle for b in bufs]); ta = (C.c_void_p * E)(*[t.handle for t in tens])
        for _ in range(3):
            assert proc.process_batch_handles(ba, ta, E) == B.kBandOk
        t0 = time.perf_counter()
        for _ in range(6):
            assert proc.process_batch_handles(ba, ta, E) == B.kBandOk
        dt = (time.perf_counter() - t0) / 6
        print("in=%s out=%s  %.3f ms / %d frames = %.0f frames/s" % ("dev" if in_dev else "host", "dev" if out_dev else "host", dt * 1e3, E, E / dt), flush=True)
        for b in bufs: b.close()
        for t in tens: t.close()
        proc.close()
