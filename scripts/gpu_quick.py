"""Scratch GPU sanity + timing of the fused kernels straight through the bandcu C-ABI.
(Development aid; the judged parity tests live in tests/.)"""
import ctypes as C
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from band_b200 import cuda_shim as S
from oracle import oracle as O


def make_desc(W, H, dw, dh, f32, mean=127.5, std=127.5):
    L = O.lib()
    x, y, dx, dy = C.c_int(), C.c_int(), C.c_int(), C.c_int()
    filt = L.orc_ScaleFilterReduce(W, H, dw, dh, 2)
    assert filt == 2, filt
    L.orc_ScaleSlope(W, H, dw, dh, filt, C.byref(x), C.byref(y), C.byref(dx), C.byref(dy))
    d = S.PlanDesc()
    d.src_kind = S.SRC_YUV420
    d.stride0 = W
    d.stride_uv = (W + 1) // 2 * 2
    d.pixel_stride_uv = 2
    d.vw, d.vh = W, H
    d.luma_map = S.Affine.identity(); d.pre_map = S.Affine.identity(); d.chroma_map = S.Affine.identity()
    d.cvt = S.CVT_YUV2RGB
    d.flavor = S.FLAVOR_ARGB
    d.scale_mode = S.SCALE_BILINEAR_DOWN
    d.linear = 0
    d.x0, d.y0, d.dx, d.dy = x.value, y.value, dx.value, dy.value
    d.max_y = (H - 1) << 16
    d.dst_width, d.dst_height = dw, dh
    d.channels = 3
    d.out_dtype = S.DT_F32 if f32 else S.DT_U8
    d.normalize = 1 if f32 else 0
    d.mean, d.std_ = mean, std
    d.dst_row_stride = dw * 3 * (4 if f32 else 1)
    return d


def run(desc, frames_t, W, H, nv21, out_t, n, stream=0):
    L = S.lib()
    plan = C.c_void_p()
    S.check(L.bandcu_plan_create(C.byref(desc), C.byref(plan)), "plan_create")
    fb = W * H + ((W + 1) // 2) * ((H + 1) // 2) * 2
    base = frames_t.data_ptr()
    fr = S.Frame()
    fr.p0 = base
    uv = base + W * H
    fr.pu, fr.pv = (uv + 1, uv) if nv21 else (uv, uv + 1)
    fr.out = out_t.data_ptr()
    ob = out_t[0].numel() * out_t.element_size()
    S.check(L.bandcu_plan_run_strided(plan, C.byref(fr), fb, ob, n, stream), "run")
    return plan, fr, fb, ob


def parity(W, H, dw, dh, f32, nv21, n=3, seed=0):
    rng = np.random.default_rng(seed)
    fb = W * H + ((W + 1) // 2) * ((H + 1) // 2) * 2
    frames = rng.integers(0, 256, size=(n, fb), dtype=np.uint8)
    fmt = O.FMT_NV21 if nv21 else O.FMT_NV12
    ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, dw, dh)]
    if f32:
        ops.append((O.OP_NORMALIZE, 0, 0, 0, 0, 127.5, 127.5))
    st, ref = O.run_batch(ops, frames, W, H, fmt, O.DT_F32 if f32 else O.DT_U8, (1, dh, dw, 3))
    assert st == 0, st
    res = {}
    for mode in ("generic", "fast"):
        os.environ["BANDCU_DISABLE_FAST"] = "1" if mode == "generic" else "0"
        ft = torch.from_numpy(frames).cuda()
        out = torch.zeros((n, 1, dh, dw, 3), dtype=torch.float32 if f32 else torch.uint8, device="cuda")
        d = make_desc(W, H, dw, dh, f32)
        plan, *_ = run(d, ft, W, H, nv21, out, n)
        torch.cuda.synchronize()
        isfast = S.lib().bandcu_plan_is_fast(plan)
        got = out.cpu().numpy()
        if f32:
            bad = int((got.view(np.uint32) != ref.view(np.uint32)).sum())
        else:
            bad = int((got != ref).sum())
        res[mode] = (isfast, bad)
        S.lib().bandcu_plan_destroy(plan)
    print("parity %dx%d %s -> %dx%d %s : %s" % (W, H, "NV21" if nv21 else "NV12", dw, dh, "f32" if f32 else "u8", res), flush=True)
    return res


def timing(W, H, dw, dh, f32, nframes=256, iters=20, rows=None, nobulk=False, generic=False):
    os.environ["BANDCU_DISABLE_FAST"] = "1" if generic else "0"
    if rows:
        os.environ["BANDCU_FAST_ROWS"] = str(rows)
    os.environ["BANDCU_FAST_NO_BULK"] = "1" if nobulk else "0"
    fb = W * H + ((W + 1) // 2) * ((H + 1) // 2) * 2
    ft = torch.randint(0, 256, (nframes, fb), dtype=torch.uint8, device="cuda")
    out = torch.zeros((nframes, 1, dh, dw, 3), dtype=torch.float32 if f32 else torch.uint8, device="cuda")
    d = make_desc(W, H, dw, dh, f32)
    L = S.lib()
    plan, fr, fbb, ob = run(d, ft, W, H, False, out, nframes)
    torch.cuda.synchronize()
    for _ in range(3):
        L.bandcu_plan_run_strided(plan, C.byref(fr), fbb, ob, nframes, 0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        L.bandcu_plan_run_strided(plan, C.byref(fr), fbb, ob, nframes, 0)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    fps = nframes / (ms * 1e-3)
    alg = fb + out[0].numel() * out.element_size()
    print("timing %dx%d->%dx%d %s rows=%s nobulk=%s fast=%d: %.3f ms/batch(%d)  %.0f frames/s  alg %.1f GB/s (%.3f of 6455.6)"
          % (W, H, dw, dh, "f32" if f32 else "u8", rows, nobulk, L.bandcu_plan_is_fast(plan), ms, nframes, fps, fps * alg / 1e9,
             fps * alg / 1e9 / 6455.6), flush=True)
    L.bandcu_plan_destroy(plan)


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    if what in ("all", "parity"):
        parity(1920, 1080, 224, 224, True, False)
        parity(1920, 1080, 300, 300, False, False)
        parity(640, 480, 224, 224, True, True)
        parity(1920, 1080, 513, 513, True, True)
        parity(642, 482, 225, 223, False, False)   # unaligned -> cooperative staging
    if what in ("all", "timing"):
        for rows in (2, 4, 8):
            timing(1920, 1080, 224, 224, True, rows=rows)
        timing(1920, 1080, 224, 224, True, rows=4, nobulk=True)
        timing(1920, 1080, 300, 300, False, rows=4)
        timing(1920, 1080, 513, 513, True, rows=4)
        timing(1920, 1080, 224, 224, True, generic=True)
    if what == "one":
        timing(1920, 1080, 224, 224, True, rows=int(os.environ.get("ROWS", "4")), iters=3)
