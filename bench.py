#!/usr/bin/env python
"""bench.py -- frames/s of the band/buffer hot path on B200 (BASELINE.json `metric`).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--frames F]

One "step" = one pass of the fused chain over one batch of F synthetic 1080p NV12 frames per GPU:
ColorSpaceConvert(kRGB) -> Resize(224,224) -> Normalize(127.5,127.5) into float32 [1,224,224,3]
tensors, driven through Band's C API (libband_c.so -> libbandcu.so). Prints ONE JSON line (rank 0).

  value      whole-job frames/s with frames and tensors resident in HBM (CUDA events on the launch stream)
  e2e        same metric through BandImageProcessorProcessBatch with HOST buffers/tensors: the pinned
             H2D upload of every frame and the D2H read of every tensor are inside the timed region
  roofline   algorithmic bytes per launch / mean launch duration of the fused kernel, vs MEASURED_PEAKS.json
  cpu_baseline  the CPU oracle (restated reference path) on this box's host cores, bounded sample (N=1 only)

--impl reference times the reference's CPU path (the oracle port; the reference itself cannot be built
here: libyuv/abseil/bazel absent) on all host threads for the same workload and prints the same line shape.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

W, H, DW, DH = 1920, 1080, 224, 224
MEAN, STD = 127.5, 127.5
FRAME_BYTES = W * H + 2 * ((W + 1) // 2) * ((H + 1) // 2)      # 3 110 400
OUT_BYTES = DW * DH * 3 * 4                                    # 602 112
ALG_BYTES = FRAME_BYTES + OUT_BYTES                            # 3 712 512 (SURVEY 8d / DESIGN.md)
WORKLOAD = "1080p NV12 -> ColorSpaceConvert(RGB) -> Resize(224,224) -> Normalize(127.5,127.5) f32 [1,224,224,3]"
METRIC = "frames/s (1080p NV12->224x224 RGB f32)"
# dram__bytes_read.sum + dram__bytes_write.sum of the fused kernel per frame, from the committed ncu --set full
# capture (profiles/r01_ncu_full_fast_v4.txt: 770.881 MB + 279.695 MB over a 512-frame launch). Below the
# algorithmic figure because the down-scaler never reads the source rows it does not sample.
NCU_TRAFFIC_PER_FRAME = (770.880512e6 + 279.695104e6) / 512


def measured_peak():
    """(GB/s, source) -- MEASURED_PEAKS.json if present, else the profiling recipe's fallback."""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def synth_frames(n_unique, seed=0x42414E44):
    rng = np.random.default_rng(seed)
    return rng.integers(0, 256, size=(n_unique, FRAME_BYTES), dtype=np.uint8)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 9:
                continue
            try:
                sm.append(float(p[1])); mx.append(float(p[2])); power.append(float(p[3]))
            except ValueError:
                continue
            for name, v in zip(names, p[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference chain on the host cores (cpu_baseline and --impl reference)
# ---------------------------------------------------------------------------------------------------
def cpu_frames_per_s(frames_per_thread, threads, repeats=1):
    """Runs the restated reference chain on `threads` host threads (one processor per thread, as the
    reference requires: buffer_processor.cc:28 is not re-entrant). Returns (frames/s, seconds)."""
    from oracle import oracle as O
    O.lib()
    ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, DW, DH), (O.OP_NORMALIZE, 0, 0, 0, 0, MEAN, STD)]
    uniq = synth_frames(min(frames_per_thread, 8))
    idx = np.arange(frames_per_thread) % uniq.shape[0]
    frames = np.ascontiguousarray(uniq[idx])

    def work():
        for _ in range(repeats):
            st, _ = O.run_batch(ops, frames, W, H, O.FMT_NV12, O.DT_F32, (1, DH, DW, 3))
            assert st == 0, st

    ths = [threading.Thread(target=work) for _ in range(threads)]
    t0 = time.perf_counter()
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    dt = time.perf_counter() - t0
    return frames_per_thread * threads * repeats / dt, dt


def host_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


# ---------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------
def dist_setup(n_gpus):
    """One process per GPU (torchrun env); NCCL carries the barrier and the timing reduction only."""
    from band_b200 import shard
    rank, local, world = shard.env_rank()
    return rank, local, world, shard.init_dist("nccl")


_ALL_CPUS = None


def unbind():
    """Back to every host CPU (the CPU baseline uses all of them)."""
    if _ALL_CPUS:
        os.sched_setaffinity(0, _ALL_CPUS)


def bind_near_gpu(local):
    """Pin this rank's threads to the CPUs of its GPU's NUMA node before any pinned allocation, so the frames and
    tensors of the e2e path sit in the memory closest to the GPU's PCIe root (one host thread per GPU).
    Returns a short description for the JSON line; never fatal."""
    if os.environ.get("BENCH_NO_AFFINITY") == "1":
        return "off"
    try:
        import torch
        bus = torch.cuda.get_device_properties(local).pci_bus_id if hasattr(torch.cuda.get_device_properties(local), "pci_bus_id") else None
        if bus is None:
            import pynvml
            pynvml.nvmlInit()
            bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(local)).busId
            if isinstance(bus, bytes):
                bus = bus.decode()
        bus = str(bus).lower()
        if len(bus.split(":")[0]) == 8:   # nvml prints an 8-digit domain, sysfs uses 4
            bus = bus[4:]
        base = "/sys/bus/pci/devices/" + bus
        node = int(open(base + "/numa_node").read().strip())
        cpus = open(base + "/local_cpulist").read().strip()
        ids = set()
        for part in cpus.split(","):
            if "-" in part:
                a, b = part.split("-")
                ids.update(range(int(a), int(b) + 1))
            elif part:
                ids.add(int(part))
        global _ALL_CPUS
        _ALL_CPUS = os.sched_getaffinity(0)
        ids &= _ALL_CPUS
        if ids:
            os.sched_setaffinity(0, ids)
        return "numa_node %d, %d cpus" % (node, len(ids))
    except Exception as e:  # noqa: BLE001
        return "unavailable (%s)" % type(e).__name__


def barrier_max(dist, local, value_ms):
    """max over ranks of a device-measured duration (no data-path collective: timing only)."""
    from band_b200 import shard
    return shard.max_over_ranks(dist, value_ms, device="cuda:%d" % local)


def sync_all(dist, local):
    import torch
    torch.cuda.synchronize(local)
    if dist is not None:
        dist.barrier()
        torch.cuda.synchronize(local)


def run_ours(args):
    import ctypes as C
    import torch
    from band_b200 import band_c as B
    from band_b200 import cuda_shim as S
    from band_b200 import hotpath

    rank, local, world, dist = dist_setup(args.gpus)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product has no CPU fallback")
    torch.cuda.set_device(local)
    affinity = bind_near_gpu(local)
    hotpath.set_device(local)
    L, SL = B.lib(), S.lib()
    ops = hotpath.chain_ops(DW, DH, True, MEAN, STD)
    F = args.frames

    # ---- resident batch: F frames + F tensors in HBM (F * 3.1 MB >> 126 MB of L2: every step streams from HBM)
    rb = hotpath.ResidentBatch(ops, F, W, H, "NV12", np.float32, (1, DH, DW, 3), device=local)
    uniq = synth_frames(16, seed=0x42414E44 + rank)
    for i in range(F):
        if L.BandCopyToDevice(rb.dev_frames + i * rb.frame_pitch, uniq[i % 16].ctypes.data, FRAME_BYTES) != B.kBandOk:
            raise SystemExit("upload failed: " + B.last_error())
    launches0 = hotpath.launch_count()
    rb.step()  # compiles the plan; creates the stream
    assert rb.is_fused, "north-star chain did not lower to the single fused launch"
    stream = rb.proc.stream
    for _ in range(max(args.warmup, 3) - 1):
        rb.step()

    def ev():
        e = C.c_void_p()
        S.check(SL.bandcu_event_create(C.byref(e)), "event_create")
        return e

    K = args.steps
    evs = [(ev(), ev()) for _ in range(K)]
    e_begin, e_end = ev(), ev()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    sync_all(dist, local)
    launches_before = hotpath.launch_count()
    S.check(SL.bandcu_event_record(e_begin, stream), "record")
    for k in range(K):
        S.check(SL.bandcu_event_record(evs[k][0], stream), "record")
        st = rb.proc.process_batch_async(rb._ba, rb._ta, F)
        S.check(SL.bandcu_event_record(evs[k][1], stream), "record")
        if st != B.kBandOk:
            raise SystemExit("ProcessBatchAsync failed: " + B.last_error())
    S.check(SL.bandcu_event_record(e_end, stream), "record")
    rb.proc.synchronize()
    sync_all(dist, local)
    clocks = sampler.stop() if rank == 0 else None
    gpu_launches = hotpath.launch_count() - launches_before
    ms = C.c_float()
    S.check(SL.bandcu_event_elapsed_ms(e_begin, e_end, C.byref(ms)), "elapsed")
    total_ms = barrier_max(dist, local, float(ms.value))
    kern_ms = []
    for a, b in evs:
        S.check(SL.bandcu_event_elapsed_ms(a, b, C.byref(ms)), "elapsed")
        kern_ms.append(float(ms.value))
    kernel_ms = float(np.mean(kern_ms))
    value = F * world * K / (total_ms * 1e-3)

    # two results of the timed batch, kept for the spot check inside the cpu_baseline leg below (the only place
    # bench.py touches the oracle on this arm)
    got_resident = rb.download([0, 1]) if rank == 0 else None
    rb.close()

    # ---- e2e: host frames (pinned) -> BandImageProcessorProcessBatch -> host tensors (pinned), per step
    E = args.e2e_frames
    host = L.BandHostAlloc(E * FRAME_BYTES)
    if not host:
        raise SystemExit("BandHostAlloc failed")
    hv = np.ctypeslib.as_array(C.cast(host, C.POINTER(C.c_uint8)), shape=(E, FRAME_BYTES))
    for i in range(E):
        hv[i] = uniq[i % 16]
    proc = B.ImageProcessor(ops)
    tens = [B.Tensor(np.float32, (1, DH, DW, 3), device=False) for _ in range(E)]
    bufs = [B.Buffer().set_raw(host + i * FRAME_BYTES, W, H, B.kBandNV12) for i in range(E)]
    ba = (C.c_void_p * E)(*[b.handle for b in bufs])
    ta = (C.c_void_p * E)(*[t.handle for t in tens])
    for _ in range(max(2, min(args.warmup, 3))):
        assert proc.process_batch_handles(ba, ta, E) == B.kBandOk, B.last_error()
    sync_all(dist, local)
    e2e_launches0 = hotpath.launch_count()
    t0 = time.perf_counter()
    checksum = 0.0
    for _ in range(args.e2e_steps):
        assert proc.process_batch_handles(ba, ta, E) == B.kBandOk, B.last_error()
        # the call is synchronous: results are in host memory; touch one value per step like a consumer would
        checksum += float(np.ctypeslib.as_array(C.cast(tens[-1].data_ptr, C.POINTER(C.c_float)), shape=(4,))[0])
    e2e_ms = (time.perf_counter() - t0) * 1e3
    e2e_ms = barrier_max(dist, local, e2e_ms)
    e2e_launches = hotpath.launch_count() - e2e_launches0
    e2e_value = E * world * args.e2e_steps / (e2e_ms * 1e-3)
    got_e2e = tens[0].numpy() if rank == 0 else None
    for b in bufs:
        b.close()
    for t in tens:
        t.close()
    proc.close()
    L.BandHostFree(host)

    peak, peak_src = measured_peak()
    achieved = ALG_BYTES * F / (kernel_ms * 1e-3) / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": max(args.warmup, 3),
        "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8 fixed-point (6/7/8-bit) -> f32 store", "data": "synthetic (seeded uniform bytes)",
        "config": {"workload": WORKLOAD, "frames_per_step_per_gpu": F, "bytes_per_frame_algorithmic": ALG_BYTES,
                   "l2_policy": "inputs larger than L2 (%.1f GB of frames per step per GPU vs 126 MB L2)" % (F * FRAME_BYTES / 1e9),
                   "parallelism": "frames sharded across %d GPU(s), no collective" % world,
                   "timing": "CUDA events on the processor's launch stream, max over ranks"},
        "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": E * FRAME_BYTES * world,
                "d2h_bytes_per_step": E * OUT_BYTES * world, "frames_per_step_per_gpu": E, "steps": args.e2e_steps,
                "timing": "host clock around synchronous BandImageProcessorProcessBatch calls (pinned host buffers and tensors), max over ranks",
                "gpu_launches": e2e_launches, "host_affinity": affinity,
                "note": "h2d/d2h_bytes_per_step count the full frames and tensors handed to the call; the fused kernel reads the "
                        "pinned frames in place over PCIe (only the source rows the resize samples cross the bus) and pushes "
                        "the tensors into pinned host memory with cp.async.bulk stores -- no staging copy either way"},
        "gpu_launches": gpu_launches,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": args.traffic if args.traffic is not None else NCU_TRAFFIC_PER_FRAME * F,
                     "traffic_source": "ncu --set full, profiles/r01_ncu_full_fast_v4.txt, scaled per frame",
                     "peak_source": peak_src, "kernel": "fast_nv_bilinear_down_kernel<float, kChromaUV, BULKOUT=false>",
                     "kernel_ms": kernel_ms, "bytes_per_launch_algorithmic": ALG_BYTES * F},
        "clocks": clocks, "parity": None,
    }
    unbind()
    if rank == 0 and world == 1 and not args.no_cpu:
        th = host_threads()
        fps1, _ = cpu_frames_per_s(4, th)                     # short probe to size the sample
        per_thread = max(4, int(fps1 / th * args.cpu_seconds))
        fps, secs = cpu_frames_per_s(per_thread, th)
        if not args.no_check:  # the same oracle, as the checker of two frames of the timed batch and one e2e frame
            from oracle import oracle as O
            o_ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, DW, DH), (O.OP_NORMALIZE, 0, 0, 0, 0, MEAN, STD)]
            st, want = O.run_batch(o_ops, uniq[:2], W, H, O.FMT_NV12, O.DT_F32, (1, DH, DW, 3))
            line["parity"] = {
                "frames_checked": 2,
                "mismatching_elements": int((got_resident.view(np.uint32) != want.view(np.uint32)).sum()),
                "e2e_mismatching_elements": int((got_e2e.view(np.uint32) != want[0].view(np.uint32)).sum())}
        line["cpu_baseline"] = {"value": fps, "unit": "frames/s", "cores": th, "kind": "port",
                                "sample": "%d frames of the same workload (%d per thread x %d threads), %.1f s; restated scalar C, no hand-written SIMD"
                                          % (per_thread * th, per_thread, th, secs)}
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        emit(line)


def run_reference(args):
    """Reference arm: the reference's own CPU path for this workload. The reference cannot be built here
    (libyuv, abseil and bazel are absent), so this is the oracle port, on all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    th = host_threads()
    fps1, _ = cpu_frames_per_s(4, th)
    per_thread = max(2, int(fps1 / th * args.ref_step_seconds))
    for _ in range(max(args.warmup, 1)):
        cpu_frames_per_s(min(per_thread, 4), th)
    t_total, n_total = 0.0, 0
    for _ in range(args.steps):
        fps, secs = cpu_frames_per_s(per_thread, th)
        t_total += secs
        n_total += per_thread * th
    value = n_total / t_total
    sample = "%d frames per step (%d per thread x %d threads) of the same workload" % (per_thread * th, per_thread, th)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": max(args.warmup, 1), "ms_per_step": t_total / args.steps * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8 fixed-point -> f32", "data": "synthetic (seeded uniform bytes)",
            "config": {"workload": WORKLOAD, "frames_per_step": per_thread * th},
            "cpu_baseline": {"value": value, "unit": "frames/s", "cores": th, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


_REAL_STDOUT = None


def guard_stdout():
    """The contract is ONE JSON line on stdout. Libraries (NCCL prints its version banner to stdout when NCCL_DEBUG is
    set) write to fd 1 behind Python's back, so fd 1 is pointed at stderr for the whole run and the line goes to a
    private duplicate of the original stdout."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    guard_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--frames", type=int, default=2048, help="frames per step per GPU (resident)")
    ap.add_argument("--e2e-frames", type=int, default=256)
    ap.add_argument("--e2e-steps", type=int, default=8)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--ref-step-seconds", type=float, default=4.0)
    ap.add_argument("--traffic", type=float, default=None, help="ncu dram bytes per launch (from profiles/), if known")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-check", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
