#!/usr/bin/env python
"""Secondary measurements: every BASELINE.json config (C1..C5) on one GPU, device-resident, through the Band C API.

  python scripts/bench_configs.py [--configs c1,c2,c3,c4,c5] [--steps K] [--cpu] > profiles/rNN_configs.jsonl

One JSON line per case: frames/s (CUDA events on the processor's launch stream), algorithmic bytes per frame
(input frame once in full + output), achieved GB/s and fraction of the measured HBM copy bandwidth,
kernel launches per step, and (with --cpu) the oracle port on ONE host thread for the same chain. bench.py stays the
driver-facing headline (north star); this script is the evidence for the other rows of DESIGN.md's table.
"""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from band_b200 import band_c as B          # noqa: E402
from band_b200 import cuda_shim as S       # noqa: E402
from band_b200 import hotpath              # noqa: E402


def peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6650.0


def event():
    e = C.c_void_p()
    S.check(S.lib().bandcu_event_create(C.byref(e)), "event")
    return e


def elapsed(a, b):
    ms = C.c_float()
    S.check(S.lib().bandcu_event_elapsed_ms(a, b, C.byref(ms)), "elapsed")
    return float(ms.value)


class ImageBatch:
    """n device frames -> n device IMAGE buffers (for outputs a tensor cannot describe: gray, YUV)."""

    def __init__(self, ops, n, W, H, fmt, out_fmt, out_w, out_h):
        L = B.lib()
        self.n = n
        self.proc = B.ImageProcessor(ops)
        self.fb = B.frame_bytes(W, H, fmt)
        self.pitch = (self.fb + 255) // 256 * 256
        self.ob = B.frame_bytes(out_w, out_h, out_fmt)
        self.opitch = (self.ob + 255) // 256 * 256
        self.din = L.BandDeviceAlloc(self.pitch * n)
        self.dout = L.BandDeviceAlloc(self.opitch * n)
        assert self.din and self.dout
        self.inb = [B.Buffer().set_raw(self.din + i * self.pitch, W, H, fmt, device=True) for i in range(n)]
        self.outb = [B.Buffer().set_raw(self.dout + i * self.opitch, out_w, out_h, out_fmt, device=True) for i in range(n)]
        self._ia = (C.c_void_p * n)(*[b.handle for b in self.inb])
        self._oa = (C.c_void_p * n)(*[b.handle for b in self.outb])

    def fill(self, uniq):
        L = B.lib()
        for i in range(self.n):
            assert L.BandCopyToDevice(self.din + i * self.pitch, uniq[i % len(uniq)].ctypes.data, self.fb) == B.kBandOk

    def step(self, async_=True):
        st = B.lib().BandImageProcessorProcessBuffers(self.proc.handle, self._ia, self._oa, self.n, 1 if async_ else 0)
        assert st == B.kBandOk, B.last_error()

    def close(self):
        for b in self.inb + self.outb:
            b.close()
        self.proc.close()
        B.lib().BandDeviceFree(self.din)
        B.lib().BandDeviceFree(self.dout)


def time_steps(step_async, sync, stream_of, steps, warmup=3):
    for _ in range(warmup):
        step_async()
    sync()
    stream = stream_of()
    e0, e1 = event(), event()
    L = S.lib()
    n0 = hotpath.launch_count()
    S.check(L.bandcu_event_record(e0, stream), "rec")
    for _ in range(steps):
        step_async()
    S.check(L.bandcu_event_record(e1, stream), "rec")
    sync()
    return elapsed(e0, e1) / steps, (hotpath.launch_count() - n0) / steps


def cpu_fps(ops_oracle, W, H, fmt_o, out_dt, out_dims, seconds=2.0):
    """Oracle port, ONE host thread, bounded sample."""
    from oracle import oracle as O
    rng = np.random.default_rng(1)
    frame = rng.integers(0, 256, size=O.frame_bytes(W, H, fmt_o), dtype=np.uint8)
    n, t0 = 0, time.perf_counter()
    while True:
        if out_dims is None:
            cur, w, h, f = frame, W, H, fmt_o
            for op in ops_oracle:
                st, cur, (w, h, f) = O.run_op(op, cur, w, h, f)
                assert st == 0
        else:
            st, _ = O.run_chain(ops_oracle, frame, W, H, fmt_o, out_dt, out_dims)
            assert st == 0, st
        n += 1
        dt = time.perf_counter() - t0
        if dt > seconds and n >= 2:
            return n / dt


NORM_O = (5, 0, 0, 0, 0, 127.5, 127.5)


def tensor_case(name, W, H, fmt, dw, dh, f32, n, steps, cpu, pre_ops=(), pre_ops_o=(), i8=False, out_dims=None):
    """chain [pre_ops] -> CSC(RGB) -> Resize -> [Normalize] into device tensors."""
    ops = list(pre_ops) + (hotpath.chain_ops(dw, dh, True, mean=128.0, std=1.0) if i8 else hotpath.chain_ops(dw, dh, f32))
    dt = np.int8 if i8 else (np.float32 if f32 else np.uint8)  # i8: Normalize(128, 1) into an int8 tensor (quantized models)
    rb = hotpath.ResidentBatch(ops, n, W, H, fmt, dt, out_dims or (1, dh, dw, 3))
    rng = np.random.default_rng(3)
    uniq = rng.integers(0, 256, size=(4, rb.frame_bytes), dtype=np.uint8)
    L = B.lib()
    for i in range(n):
        assert L.BandCopyToDevice(rb.dev_frames + i * rb.frame_pitch, uniq[i % 4].ctypes.data, rb.frame_bytes) == B.kBandOk
    rb.step()
    ms, launches = time_steps(lambda: rb.proc.process_batch_async(rb._ba, rb._ta, n), rb.proc.synchronize,
                              lambda: rb.proc.stream, steps)
    alg = rb.frame_bytes + rb.out_bytes
    fps = n / (ms * 1e-3)
    line = {"case": name, "frames_per_s": fps, "ms_per_step": ms, "frames_per_step": n, "launches_per_step": launches,
            "fused_single_launch": rb.is_fused, "alg_bytes_per_frame": alg, "achieved_GBps": fps * alg / 1e9,
            "frac_of_measured_hbm": fps * alg / 1e9 / peak()}
    if cpu:
        from oracle import oracle as O
        o_ops = list(pre_ops_o) + [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, dw, dh)] + ([NORM_O] if f32 else [])
        line["cpu_oracle_1thread_fps"] = cpu_fps(o_ops, W, H, getattr(O, "FMT_" + fmt), O.DT_F32 if f32 else O.DT_U8,
                                                 (1, dh, dw, 3))
    rb.close()
    return line


def c1_latency(steps):
    """Single 640x480 NV21 frame -> 224x224 f32: latency of one synchronous BandImageProcessorProcess call."""
    W, H = 640, 480
    ops = hotpath.chain_ops(224, 224, True)
    out = []
    for resident in (True, False):
        proc = B.ImageProcessor(ops)
        t = B.Tensor(np.float32, (1, 224, 224, 3), device=resident)
        fb = B.frame_bytes(W, H, B.kBandNV21)
        rng = np.random.default_rng(5)
        frame = rng.integers(0, 256, size=fb, dtype=np.uint8)
        L = B.lib()
        if resident:
            dev = L.BandDeviceAlloc(fb)
            L.BandCopyToDevice(dev, frame.ctypes.data, fb)
            buf = B.Buffer().set_raw(dev, W, H, B.kBandNV21, device=True)
        else:
            host = L.BandHostAlloc(fb)
            C.memmove(host, frame.ctypes.data, fb)
            buf = B.Buffer().set_raw(host, W, H, B.kBandNV21)
        for _ in range(20):
            assert proc.process(buf, t) == B.kBandOk
        ts = []
        for _ in range(max(steps, 200)):
            t0 = time.perf_counter()
            proc.process(buf, t)
            ts.append((time.perf_counter() - t0) * 1e6)
        ts = np.array(ts)
        out.append({"case": "c1_single_640x480_nv21_to_224_f32_" + ("resident" if resident else "pinned_host"),
                    "latency_us_median": float(np.median(ts)), "latency_us_p10": float(np.percentile(ts, 10)),
                    "latency_us_p90": float(np.percentile(ts, 90)), "note": "host clock around the synchronous call"})
        # the same call recorded into a CUDA graph (BandImageProcessorRecordBatch) and replayed
        rec = proc.record_batch([buf], [t])
        for _ in range(20):
            assert proc.replay(rec) == B.kBandOk
        ts = []
        for _ in range(max(steps, 200)):
            t0 = time.perf_counter()
            proc.replay(rec)
            ts.append((time.perf_counter() - t0) * 1e6)
        ts = np.array(ts)
        out.append({"case": "c1_single_640x480_nv21_to_224_f32_" + ("resident" if resident else "pinned_host") + "_cuda_graph",
                    "latency_us_median": float(np.median(ts)), "latency_us_p10": float(np.percentile(ts, 10)),
                    "latency_us_p90": float(np.percentile(ts, 90)), "note": "host clock around a synchronous graph replay"})
        proc.release(rec)
        buf.close(); t.close(); proc.close()
    return out


C5_PAIRS = [(2048, 64), (2048, 128), (2048, 256), (2048, 512), (2048, 1024), (64, 128), (64, 256), (64, 512),
            (64, 1024), (64, 2048)]


def c5_cases(steps, cpu, pairs=None):
    lines = []
    from oracle import oracle as O
    for S_, D_ in (pairs or C5_PAIRS):
        n = int(max(8, min(2048, 1.0e9 // (4 * S_ * S_ + 3 * D_ * D_))))
        # (i) AutoConvert (empty builder) into a u8 RGB tensor
        rb = hotpath.ResidentBatch([], n, S_, S_, "RGBA", np.uint8, (1, D_, D_, 3))
        rng = np.random.default_rng(7)
        uniq = rng.integers(0, 256, size=(2, rb.frame_bytes), dtype=np.uint8)
        for i in range(n):
            B.lib().BandCopyToDevice(rb.dev_frames + i * rb.frame_pitch, uniq[i % 2].ctypes.data, rb.frame_bytes)
        rb.step()
        ms, launches = time_steps(lambda: rb.proc.process_batch_async(rb._ba, rb._ta, n), rb.proc.synchronize,
                                  lambda: rb.proc.stream, steps)
        alg = 4 * S_ * S_ + 3 * D_ * D_
        fps = n / (ms * 1e-3)
        line = {"case": "c5_rgba_%d_to_rgb_%d" % (S_, D_), "frames_per_s": fps, "ms_per_step": ms, "frames_per_step": n,
                "launches_per_step": launches, "alg_bytes_per_frame": alg, "achieved_GBps": fps * alg / 1e9,
                "frac_of_measured_hbm": fps * alg / 1e9 / peak()}
        if cpu:
            line["cpu_oracle_1thread_fps"] = cpu_fps([(O.OP_AUTO,)], S_, S_, O.FMT_RGBA, O.DT_U8, (1, D_, D_, 3), 1.0)
        lines.append(line)
        rb.close()
        # (ii) CSC(Gray) -> Resize into a gray image buffer
        ops = [(B.BAND_COLOR_SPACE_CONVERT, B.kBandGrayScale), (B.BAND_RESIZE, D_, D_)]
        ib = ImageBatch(ops, n, S_, S_, B.kBandRGBA, B.kBandGrayScale, D_, D_)
        ib.fill(uniq)
        ib.step(False)
        ms, launches = time_steps(ib.step, ib.proc.synchronize, lambda: ib.proc.stream, steps)
        alg = 4 * S_ * S_ + D_ * D_
        fps = n / (ms * 1e-3)
        line = {"case": "c5_rgba_%d_to_gray_%d" % (S_, D_), "frames_per_s": fps, "ms_per_step": ms, "frames_per_step": n,
                "launches_per_step": launches, "alg_bytes_per_frame": alg, "achieved_GBps": fps * alg / 1e9,
                "frac_of_measured_hbm": fps * alg / 1e9 / peak()}
        if cpu:
            line["cpu_oracle_1thread_fps"] = cpu_fps([(O.OP_CSC, O.FMT_GRAY), (O.OP_RESIZE, D_, D_)], S_, S_, O.FMT_RGBA,
                                                     None, None, 1.0)
        lines.append(line)
        ib.close()
    return lines


def image_case(name, ops, W, H, fmt, out_fmt, ow, oh, n, steps):
    """Single-operator image -> image chains (the rows of SURVEY 8a that do not end in a tensor)."""
    ib = ImageBatch(ops, n, W, H, fmt, out_fmt, ow, oh)
    rng = np.random.default_rng(11)
    uniq = rng.integers(0, 256, size=(2, ib.fb), dtype=np.uint8)
    ib.fill(uniq)
    ib.step(False)
    ms, launches = time_steps(ib.step, ib.proc.synchronize, lambda: ib.proc.stream, steps)
    alg = ib.fb + ib.ob
    fps = n / (ms * 1e-3)
    ib.close()
    return {"case": name, "frames_per_s": fps, "ms_per_step": ms, "frames_per_step": n, "launches_per_step": launches,
            "alg_bytes_per_frame": alg, "achieved_GBps": fps * alg / 1e9, "frac_of_measured_hbm": fps * alg / 1e9 / peak()}


def r_cases(steps):
    W, H, n = 1920, 1080, 256
    return [
        image_case("g_1080p_nv21_resize_640x360", [(B.BAND_RESIZE, 640, 360)], W, H, B.kBandNV21, B.kBandNV21, 640, 360, n, steps),
        # YUV-domain Resize (ResizeNv / ResizeYv, libyuv_image_operator.cc:376-435) at ratios that take libyuv's other paths:
        # 1/3 above is point-sampled; 2/3 bilinear down, 1/2 the 2x2 box, 0.55 a free bilinear ratio, 2x the up-samplers
        image_case("g_1080p_nv21_resize_1280x720", [(B.BAND_RESIZE, 1280, 720)], W, H, B.kBandNV21, B.kBandNV21, 1280, 720, n, steps),
        image_case("g_1080p_yv12_resize_1280x720", [(B.BAND_RESIZE, 1280, 720)], W, H, B.kBandYV12, B.kBandYV12, 1280, 720, n, steps),
        image_case("g_1080p_nv21_resize_960x540", [(B.BAND_RESIZE, 960, 540)], W, H, B.kBandNV21, B.kBandNV21, 960, 540, n, steps),
        image_case("g_1080p_nv21_resize_1056x594", [(B.BAND_RESIZE, 1056, 594)], W, H, B.kBandNV21, B.kBandNV21, 1056, 594, n, steps),
        image_case("g_1080p_yv12_resize_416x234", [(B.BAND_RESIZE, 416, 234)], W, H, B.kBandYV12, B.kBandYV12, 416, 234, n, steps),
        image_case("g_540p_nv21_resize_up_1920x1080", [(B.BAND_RESIZE, 1920, 1080)], 960, 540, B.kBandNV21, B.kBandNV21, 1920, 1080, n, steps),
    ]


def g_cases(steps):
    W, H, n = 1920, 1080, 256
    csc = B.BAND_COLOR_SPACE_CONVERT
    return [
        image_case("g_1080p_nv21_to_rgb", [(csc, B.kBandRGB)], W, H, B.kBandNV21, B.kBandRGB, W, H, n, steps),
        image_case("g_1080p_nv21_to_rgba", [(csc, B.kBandRGBA)], W, H, B.kBandNV21, B.kBandRGBA, W, H, n, steps),
        image_case("g_1080p_yv12_to_rgb", [(csc, B.kBandRGB)], W, H, B.kBandYV12, B.kBandRGB, W, H, n, steps),
        image_case("g_1080p_nv21_to_gray", [(csc, B.kBandGrayScale)], W, H, B.kBandNV21, B.kBandGrayScale, W, H, n, steps),
        image_case("g_1080p_rgb_to_gray", [(csc, B.kBandGrayScale)], W, H, B.kBandRGB, B.kBandGrayScale, W, H, n, steps),
        image_case("g_1080p_rgba_to_nv21", [(csc, B.kBandNV21)], W, H, B.kBandRGBA, B.kBandNV21, W, H, n, steps),
        image_case("g_1080p_rgb_to_yv12", [(csc, B.kBandYV12)], W, H, B.kBandRGB, B.kBandYV12, W, H, n, steps),
        image_case("g_1080p_nv12_to_yv12_repack", [(csc, B.kBandYV12)], W, H, B.kBandNV12, B.kBandYV12, W, H, n, steps),
        image_case("g_1080p_yv12_to_nv21_repack", [(csc, B.kBandNV21)], W, H, B.kBandYV12, B.kBandNV21, W, H, n, steps),
        image_case("g_1080p_yv12_rot90", [(B.BAND_ROTATE, 90)], W, H, B.kBandYV12, B.kBandYV12, H, W, n, steps),
        image_case("g_1080p_rgba_rot270", [(B.BAND_ROTATE, 270)], W, H, B.kBandRGBA, B.kBandRGBA, H, W, n, steps),
        image_case("g_1080p_gray_rot90", [(B.BAND_ROTATE, 90)], W, H, B.kBandGrayScale, B.kBandGrayScale, H, W, n, steps),
        image_case("g_1080p_nv21_rot90", [(B.BAND_ROTATE, 90)], W, H, B.kBandNV21, B.kBandNV21, H, W, n, steps),
        image_case("g_1080p_rgb_rot90", [(B.BAND_ROTATE, 90)], W, H, B.kBandRGB, B.kBandRGB, H, W, n, steps),
        image_case("g_1080p_nv21_flip_h", [(B.BAND_FLIP, 1, 0)], W, H, B.kBandNV21, B.kBandNV21, W, H, n, steps),
        image_case("g_1080p_nv21_crop_center_720p", [(B.BAND_CROP, 320, 180, 1599, 899)], W, H, B.kBandNV21, B.kBandNV21,
                   1280, 720, n, steps),
        image_case("g_1080p_nv21_resize_640x360", [(B.BAND_RESIZE, 640, 360)], W, H, B.kBandNV21, B.kBandNV21, 640, 360, n, steps),
        image_case("g_1080p_rgb_resize_640x360", [(B.BAND_RESIZE, 640, 360)], W, H, B.kBandRGB, B.kBandRGB, 640, 360, n, steps),
    ]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="c1,c2,c3,c4,c5,x,g,r")
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--cpu", action="store_true")
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--c5-pairs", default="", help="e.g. 64:2048,2048:1024 (default: the full sweep)")
    a = ap.parse_args()
    hotpath.set_device(a.device)
    want = set(a.configs.split(","))
    out = []
    if "ns" in want or "c4" in want:
        out.append(tensor_case("north_star_1080p_nv12_to_224_f32", 1920, 1080, "NV12", 224, 224, True, 1024, a.steps, a.cpu))
    if "c1" in want:
        out += c1_latency(a.steps)
    if "c1" in want or "c1b" in want:
        out.append(tensor_case("c1_batch_640x480_nv21_to_224_f32", 640, 480, "NV21", 224, 224, True, 4096, a.steps, a.cpu))
    if "c2" in want:
        out.append(tensor_case("c2_1080p_nv12_to_300_u8", 1920, 1080, "NV12", 300, 300, False, 1024, a.steps, a.cpu))
    if "c3" in want:
        from oracle import oracle as O
        out.append(tensor_case("c3_4k_yv12_rot90_crop_to_512_f32", 3840, 2160, "YV12", 512, 512, True, 128, a.steps, a.cpu,
                               pre_ops=[(B.BAND_ROTATE, 90), (B.BAND_CROP, 0, 840, 2159, 2999)],
                               pre_ops_o=[(O.OP_ROTATE, 90), (O.OP_CROP, 0, 840, 2159, 2999)]))
    if "c4" in want:
        parts = [tensor_case("c4_1080p_nv21_to_224_f32", 1920, 1080, "NV21", 224, 224, True, 683, a.steps, a.cpu),
                 tensor_case("c4_1080p_nv21_to_300_u8", 1920, 1080, "NV21", 300, 300, False, 683, a.steps, a.cpu),
                 tensor_case("c4_1080p_nv21_to_513_f32", 1920, 1080, "NV21", 513, 513, True, 682, a.steps, a.cpu)]
        out += parts
        tot_frames = sum(p["frames_per_step"] for p in parts)
        tot_ms = sum(p["ms_per_step"] for p in parts)
        tot_bytes = sum(p["frames_per_step"] * p["alg_bytes_per_frame"] for p in parts)
        out.append({"case": "c4_mixed_2048_frames_per_gpu (224 f32 / 300 u8 / 513 f32 round-robin, three launches)",
                    "frames_per_s": tot_frames / (tot_ms * 1e-3), "ms_per_step": tot_ms, "frames_per_step": tot_frames,
                    "achieved_GBps": tot_bytes / (tot_ms * 1e-3) / 1e9,
                    "frac_of_measured_hbm": tot_bytes / (tot_ms * 1e-3) / 1e9 / peak()})
    if "x" in want:  # beyond BASELINE: planar 1080p, and the portrait-camera case (NV21 sensor frame + Rotate 90)
        from oracle import oracle as O
        out.append(tensor_case("x_1080p_yv12_to_224_f32", 1920, 1080, "YV12", 224, 224, True, 1024, a.steps, a.cpu))
        out.append(tensor_case("x_1080p_nv21_rot90_to_224_f32", 1920, 1080, "NV21", 224, 224, True, 1024, a.steps, a.cpu,
                               pre_ops=[(B.BAND_ROTATE, 90)], pre_ops_o=[(O.OP_ROTATE, 90)]))
        out.append(tensor_case("x_1080p_yv12_rot270_to_300_u8", 1920, 1080, "YV12", 300, 300, False, 1024, a.steps, a.cpu,
                               pre_ops=[(B.BAND_ROTATE, 270)], pre_ops_o=[(O.OP_ROTATE, 270)]))
        out.append(tensor_case("x_1080p_nv21_to_224_i8_normalized", 1920, 1080, "NV21", 224, 224, False, 1024, a.steps, False,
                               i8=True))
        out.append(tensor_case("x_1080p_nv12_to_rgb_f32_no_resize", 1920, 1080, "NV12", 0, 0, True, 64, a.steps, False, out_dims=(1, 1080, 1920, 3)))
        out.append(tensor_case("x_640x480_nv21_up_to_1280x960_u8", 640, 480, "NV21", 1280, 960, False, 256, a.steps, False))
        out.append(tensor_case("x_320x240_yv12_up_to_1024x768_f32", 320, 240, "YV12", 1024, 768, True, 128, a.steps, False))
        out.append(tensor_case("x_1080p_nv21_rot180_to_224_f32", 1920, 1080, "NV21", 224, 224, True, 1024, a.steps, a.cpu,
                               pre_ops=[(B.BAND_ROTATE, 180)], pre_ops_o=[(O.OP_ROTATE, 180)]))
    if "r" in want:  # Resize in the YUV domain, every libyuv path
        for r in r_cases(a.steps):
            print(json.dumps(r), flush=True)
    if "g" in want:  # single-operator image -> image chains
        out += g_cases(a.steps)
    if "c5" in want:
        pairs = [tuple(int(v) for v in p.split(":")) for p in a.c5_pairs.split(",") if p] or None
        out += c5_cases(a.steps, a.cpu, pairs)
    for line in out:
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
