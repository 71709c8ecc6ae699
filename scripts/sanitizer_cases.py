#!/usr/bin/env python
"""A trimmed pass over EVERY kernel family of libbandcu.so, sized for compute-sanitizer (memcheck / racecheck /
synccheck run the kernels 10-100x slower): small frames, odd widths, unaligned pointers, both staging modes of the
row-staged kernel (cp.async.bulk and producer-warp byte copies) and its BULKOUT variant. Every case is also compared
with the CPU oracle, so a sanitizer run doubles as a parity run.

  compute-sanitizer --tool memcheck  python scripts/sanitizer_cases.py > profiles/rNN_sanitizer_memcheck.log
  compute-sanitizer --tool racecheck python scripts/sanitizer_cases.py
  compute-sanitizer --tool synccheck python scripts/sanitizer_cases.py
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from band_b200 import band_c as B      # noqa: E402
from oracle import oracle as O         # noqa: E402
import gpu_util as G                   # noqa: E402

NORM = (5, 0, 0, 0, 0, 127.5, 127.5)
os.environ["BANDCU_TRACE"] = "1"
fails = []


def check(name, got, want):
    ok = got is not None and want is not None and got.shape == want.shape and \
        np.array_equal(np.ascontiguousarray(got).view(np.uint8), np.ascontiguousarray(want).view(np.uint8))
    print("%-72s %s" % (name, "ok" if ok else "MISMATCH"), flush=True)
    if not ok:
        fails.append(name)


def tensor_case(name, ops, W, H, fmt, out_dt, dims, n=2, resident=False, env=None):
    for k, v in (env or {}).items():
        os.environ[k] = v
    try:
        frames = G.rand_frames(O, n, W, H, fmt, seed=len(name))
        st, want = G.oracle_chain(O, ops, frames, W, H, fmt, out_dt, dims)
        assert st == 0, (name, st)
        st, got = G.gpu_chain(ops, frames, W, H, fmt, out_dt, dims, device_resident=resident)
        assert st == 0, (name, B.last_error())
        check(name, got, want)
    finally:
        for k in (env or {}):
            del os.environ[k]


def pinned_case(name, ops, W, H, fmt, dims, n=3, misalign=0):
    """pinned host frames and tensors: in-place PCIe reads + BULKOUT stores (misalign != 0: the fallbacks)."""
    L = B.lib()
    frames = G.rand_frames(O, n, W, H, fmt, seed=7 + misalign)
    st, want = G.oracle_chain(O, ops, frames, W, H, fmt, O.DT_F32, dims)
    assert st == 0
    fb = frames.shape[1]
    pitch = (fb + 255) // 256 * 256
    host = L.BandHostAlloc(pitch * n + 64)
    bufs = []
    for i in range(n):
        p = host + misalign + i * pitch
        C.memmove(p, frames[i].ctypes.data, fb)
        bufs.append(B.Buffer().set_raw(p, W, H, fmt))
    tens = [B.Tensor(np.float32, dims, device=False) for _ in range(n)]
    proc = B.ImageProcessor(G.to_capi_ops(ops))
    assert proc.process_batch(bufs, tens) == B.kBandOk, B.last_error()
    check(name, np.stack([t.numpy() for t in tens]), want)
    for x in bufs + tens + [proc]:
        x.close()
    L.BandHostFree(host)


def op_case(name, op, W, H, fmt):
    frame = G.rand_frames(O, 1, W, H, fmt, seed=len(name))[0]
    st, want, info = O.run_op(op, frame, W, H, fmt)
    assert st == 0, name
    st, got, ginfo = G.gpu_op(op, frame, W, H, fmt)
    assert st == 0, (name, B.last_error())
    check(name, got, want)


CSC_RGB = (O.OP_CSC, O.FMT_RGB)
# ---- row-staged kernel (bandcu_fast.cu)
tensor_case("row-staged NV12 640x480 -> 224 f32, bulk staging, resident", [CSC_RGB, (O.OP_RESIZE, 224, 224), NORM], 640, 480,
            O.FMT_NV12, O.DT_F32, (1, 224, 224, 3), resident=True)
tensor_case("row-staged NV21 640x480 -> 300 u8, BANDCU_FAST_NO_BULK=1", [CSC_RGB, (O.OP_RESIZE, 300, 300)], 640, 480,
            O.FMT_NV21, O.DT_U8, (1, 300, 300, 3), resident=True, env={"BANDCU_FAST_NO_BULK": "1"})
tensor_case("row-staged NV12 641x479 (odd dims: byte staging) -> 224 f32", [CSC_RGB, (O.OP_RESIZE, 224, 224), NORM], 641, 479,
            O.FMT_NV12, O.DT_F32, (1, 224, 224, 3), resident=True)
tensor_case("row-staged YV12 640x480 planar chroma -> 224 f32", [CSC_RGB, (O.OP_RESIZE, 224, 224), NORM], 640, 480,
            O.FMT_YV12, O.DT_F32, (1, 224, 224, 3), resident=True)
tensor_case("row-staged NV21 crop -> 160x120 f32 (crop folded into plane offsets)",
            [(O.OP_CROP, 32, 16, 543, 399), CSC_RGB, (O.OP_RESIZE, 160, 120), NORM], 640, 480, O.FMT_NV21, O.DT_F32, (1, 120, 160, 3),
            resident=True)
tensor_case("row-staged NV21 640x480 -> 224 int8 (Normalize 128,1 byte table)",
            [CSC_RGB, (O.OP_RESIZE, 224, 224), (5, 0, 0, 0, 0, 128.0, 1.0)], 640, 480, O.FMT_NV21, O.DT_I8, (1, 224, 224, 3), resident=True)
pinned_case("row-staged NV12 pinned host in/out: in-place reads + BULKOUT stores", [CSC_RGB, (O.OP_RESIZE, 224, 224), NORM], 640, 480,
            O.FMT_NV12, (1, 224, 224, 3))
pinned_case("row-staged NV12 pinned host frames at +4 bytes (unaligned: byte staging, direct stores)",
            [CSC_RGB, (O.OP_RESIZE, 224, 224), NORM], 640, 480, O.FMT_NV12, (1, 224, 224, 3), misalign=4)
tensor_case("staged pipeline: pageable host frames (3 streams, chunks)", [CSC_RGB, (O.OP_RESIZE, 224, 224), NORM], 640, 480,
            O.FMT_NV12, O.DT_F32, (1, 224, 224, 3), n=5, env={"BAND_CHUNK_FRAMES": "2"})
# ---- tile-staged kernel (bandcu_tile.cu)
tensor_case("tile-staged NV21 Rotate 90 -> 224 f32", [(O.OP_ROTATE, 90), CSC_RGB, (O.OP_RESIZE, 224, 224), NORM], 640, 480,
            O.FMT_NV21, O.DT_F32, (1, 224, 224, 3), resident=True)
tensor_case("tile-staged YV12 Rotate 270 -> 150x200 u8", [(O.OP_ROTATE, 270), CSC_RGB, (O.OP_RESIZE, 150, 200)], 640, 480,
            O.FMT_YV12, O.DT_U8, (1, 200, 150, 3), resident=True)
tensor_case("tile-staged NV12 Flip h+v, ratio 4 (per-sample staging) -> 160x120 f32", [(O.OP_FLIP, 1, 1), CSC_RGB, (O.OP_RESIZE, 160, 120), NORM],
            1280, 960, O.FMT_NV12, O.DT_F32, (1, 120, 160, 3), n=1, resident=True)
tensor_case("tile-staged YV12 Rotate 90 + Crop -> 128 f32 (C3 shape, small)",
            [(O.OP_ROTATE, 90), (O.OP_CROP, 0, 80, 479, 559), CSC_RGB, (O.OP_RESIZE, 128, 128), NORM], 640, 480, O.FMT_YV12,
            O.DT_F32, (1, 128, 128, 3), resident=True)
# ---- packed kernels (bandcu_packed.cu) and generic (bandcu_generic.cu)
tensor_case("packed RGBA 256 -> RGB 160 (bilinear down <= 2, AutoConvert)", [(O.OP_AUTO,)], 256, 256, O.FMT_RGBA, O.DT_U8, (1, 160, 160, 3), resident=True)
tensor_case("packed RGBA 256 -> RGB 128 (2x box)", [(O.OP_AUTO,)], 256, 256, O.FMT_RGBA, O.DT_U8, (1, 128, 128, 3), resident=True)
tensor_case("packed RGBA 64 -> RGB 200 (row-streaming up-scale, word interleaved)", [(O.OP_AUTO,)], 64, 64, O.FMT_RGBA, O.DT_U8, (1, 200, 200, 3), resident=True)
tensor_case("packed RGBA 64 -> RGB 512 (row-streaming up-scale, 128-bit stores)", [(O.OP_AUTO,)], 64, 64, O.FMT_RGBA, O.DT_U8, (1, 512, 512, 3), resident=True)
tensor_case("packed RGBA 64 -> RGB 128 f32 (AutoConvert + DataTypeConvert)", [(O.OP_AUTO,)], 64, 64, O.FMT_RGBA, O.DT_F32, (1, 128, 128, 3), resident=True)
tensor_case("generic RGBA 512 -> RGB 20 (ratio 25.6)", [(O.OP_AUTO,)], 512, 512, O.FMT_RGBA, O.DT_U8, (1, 20, 20, 3), resident=True)
tensor_case("generic NV21 96x64 -> 224 f32 (4:2:0 up-scale)", [CSC_RGB, (O.OP_RESIZE, 224, 224), NORM], 96, 64, O.FMT_NV21, O.DT_F32,
            (1, 224, 224, 3), resident=True)
tensor_case("generic NV21 321x241 Rotate 90 -> 100x150 u8 (odd dims behind an index map)", [(O.OP_ROTATE, 90), CSC_RGB, (O.OP_RESIZE, 100, 150)],
            321, 241, O.FMT_NV21, O.DT_U8, (1, 150, 100, 3), resident=True)
tensor_case("Normalize without Resize: NV12 320x240 -> RGB f32", [CSC_RGB, NORM], 320, 240, O.FMT_NV12, O.DT_F32, (1, 240, 320, 3), resident=True)
# ---- same-size conversions (bandcu_copy.cu) and plane kernels, as stand-alone operators
op_case("copy-rows NV21 640x480 -> RGB", (O.OP_CSC, O.FMT_RGB), 640, 480, O.FMT_NV21)
op_case("copy-rows NV12 644x480 -> RGBA (4-pixel form)", (O.OP_CSC, O.FMT_RGBA), 644, 480, O.FMT_NV12)
op_case("copy-rows YV12 640x480 -> RGB", (O.OP_CSC, O.FMT_RGB), 640, 480, O.FMT_YV12)
op_case("packed copy RGBA 640x480 -> RGB", (O.OP_CSC, O.FMT_RGB), 640, 480, O.FMT_RGBA)
op_case("packed copy RGB 640x480 -> gray", (O.OP_CSC, O.FMT_GRAY), 640, 480, O.FMT_RGB)
op_case("packed copy RGBA 322x241 -> gray (pixel-per-thread fallback)", (O.OP_CSC, O.FMT_GRAY), 322, 241, O.FMT_RGBA)
op_case("luma copy NV21 640x480 -> gray", (O.OP_CSC, O.FMT_GRAY), 640, 480, O.FMT_NV21)
op_case("rgb_to_yuv420 RGBA 640x480 -> NV21", (O.OP_CSC, O.FMT_NV21), 640, 480, O.FMT_RGBA)
op_case("rgb_to_yuv420 RGB 321x241 -> YV12 (odd dims)", (O.OP_CSC, O.FMT_YV12), 321, 241, O.FMT_RGB)
op_case("yuv_map NV12 640x480 -> YV12 (chroma repack)", (O.OP_CSC, O.FMT_YV12), 640, 480, O.FMT_NV12)
op_case("plane_turn NV21 640x480 Rotate 90", (O.OP_ROTATE, 90), 640, 480, O.FMT_NV21)
op_case("plane_turn YV12 642x482 Rotate 270 (word-unaligned)", (O.OP_ROTATE, 270), 642, 482, O.FMT_YV12)
op_case("plane_map_tile RGB 640x480 Rotate 90", (O.OP_ROTATE, 90), 640, 480, O.FMT_RGB)
op_case("plane_map_tile RGBA 321x241 Rotate 270", (O.OP_ROTATE, 270), 321, 241, O.FMT_RGBA)
op_case("plane_rows16 NV21 640x480 Flip horizontal", (O.OP_FLIP, 1, 0), 640, 480, O.FMT_NV21)
op_case("plane_rows16 NV12 640x480 Crop (odd origin)", (O.OP_CROP, 33, 17, 544, 400), 640, 480, O.FMT_NV12)
op_case("plane map gray 641x479 Rotate 180", (O.OP_ROTATE, 180), 641, 479, O.FMT_GRAY)
op_case("YUV-domain Resize NV21 640x480 -> 320x200", (O.OP_RESIZE, 320, 200), 640, 480, O.FMT_NV21)
op_case("YUV-domain Resize YV12 640x480 -> 1000x700 (up)", (O.OP_RESIZE, 1000, 700), 640, 480, O.FMT_YV12)
op_case("gray Resize 640x480 -> 480x360 (Down34)", (O.OP_RESIZE, 480, 360), 640, 480, O.FMT_GRAY)
op_case("gray Resize 640x480 -> 240x180 (Down38)", (O.OP_RESIZE, 240, 180), 640, 480, O.FMT_GRAY)
op_case("RGB Resize 640x480 -> 333x222", (O.OP_RESIZE, 333, 222), 640, 480, O.FMT_RGB)

print("sanitizer_cases: %d case(s) differ from the oracle" % len(fails), flush=True)
sys.exit(1 if fails else 0)
