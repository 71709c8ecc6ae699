#!/usr/bin/env python
"""bench.py -- frames/s of the band/buffer hot path on B200 (BASELINE.json `metric`).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--frames F]

One "step" = one pass of the fused chain over one batch of F synthetic 1080p NV12 frames per GPU:
ColorSpaceConvert(kRGB) -> Resize(224,224) -> Normalize(127.5,127.5) into float32 [1,224,224,3]
tensors, driven through Band's C API (libband_c.so -> libbandcu.so). Prints ONE JSON line (rank 0).

  value         whole-job frames/s with frames and tensors resident in HBM (CUDA events on the launch stream)
  sustained     the same step back to back for >= 2 s (outside the K driver-timed steps), with SM clocks under load
  e2e           same metric through BandImageProcessorProcessBatch with HOST buffers/tensors (pinned): the host->device
                read of every frame and the device->host write of every tensor are inside the timed region
  e2e_pageable  the unmodified caller: frames in plain malloc'ed memory (staged pipeline, and registration cache)
  roofline      frac = NECESSARY bytes (band_b200/traffic_model.py: sampled rows x touched sectors + output) per launch
                / mean launch duration / MEASURED_PEAKS.json; frac_algorithmic = the SURVEY 8d figure (whole frame +
                output), which exceeds 1 because unsampled rows are never read; traffic = ncu DRAM bytes per launch read
                from the committed capture named in traffic_source
  secondary     the other BASELINE configs (C1 latency, C2, C3, C4-mixed through the sharded API, C5 extremes)
  cpu_baseline  the CPU oracle (restated reference path) on this box's host cores, bounded sample (N=1 only)

--impl reference times the reference's CPU path (the oracle port; the reference itself cannot be built
here: libyuv/abseil/bazel absent) on all host threads for the same workload and prints the same line shape.
"""
import argparse
import hashlib
import json
import os
import re
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
KERNEL = "fast_nv_bilinear_down_kernel<float, kChromaUV, BULKOUT=false>"
# ncu --set full summaries of the fused kernel (scripts/ncu_summary.py output), newest first. The first one present is
# used; each starts with a "# frames_per_launch: F" line (the r01 capture predates that header: 512 frames).
NCU_CAPTURES = [("profiles/r02_ncu_full_fast_F2048.txt", None), ("profiles/r01_ncu_full_fast_v4.txt", 512)]
NCU_KERNEL_TAG = "fast_nv_bilinear_down_kernel<float, 0, 0>"


def measured_peak():
    """(GB/s, source) -- MEASURED_PEAKS.json if present, else the profiling recipe's fallback."""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def git_blob_sha(path):
    """git's blob id of a file, computed here (no git binary needed on the GPU box)."""
    data = open(path, "rb").read()
    return hashlib.sha1(b"blob %d\0" % len(data) + data).hexdigest()


_UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def load_ncu_capture():
    """Per-FRAME DRAM traffic and pipe utilisation of the fused kernel from the committed ncu summary. Returns a dict
    (bytes_per_frame, issue_active_pct, alu_pipe_pct, file, git_blob, frames_per_launch, launches) or None."""
    for rel, frames in NCU_CAPTURES:
        path = os.path.join(ROOT, rel)
        if not os.path.exists(path):
            continue
        rd, wr, issue, alu, dur = [], [], [], [], []
        take = False
        for ln in open(path):
            m = re.match(r"#\s*frames_per_launch:\s*(\d+)", ln)
            if m:
                frames = int(m.group(1))
            if ln.startswith("== launch"):
                take = NCU_KERNEL_TAG in ln
                continue
            if not take:
                continue
            p = ln.split()
            if len(p) < 2:
                continue
            try:
                val = float(p[1])
            except ValueError:
                continue
            unit = p[2] if len(p) > 2 else ""
            if p[0] == "dram__bytes_read.sum":
                rd.append(val * _UNIT.get(unit, 1.0))
            elif p[0] == "dram__bytes_write.sum":
                wr.append(val * _UNIT.get(unit, 1.0))
            elif p[0] == "smsp__issue_active.avg.pct_of_peak_sustained_active":
                issue.append(val)
            elif p[0] == "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active":
                alu.append(val)
            elif p[0] == "gpu__time_duration.sum":
                dur.append(val)
        if not rd or not wr or not frames:
            continue
        return {"bytes_per_frame": float(np.mean(rd) + np.mean(wr)) / frames, "issue_active_pct": float(np.mean(issue)) if issue else None,
                "alu_pipe_pct": float(np.mean(alu)) if alu else None, "file": rel, "git_blob": git_blob_sha(path),
                "frames_per_launch": frames, "launches": len(rd), "duration_us_under_ncu": float(np.mean(dur)) if dur else None}
    return None


def synth_frames(n_unique, seed=0x42414E44):
    rng = np.random.default_rng(seed)
    return rng.integers(0, 256, size=(n_unique, FRAME_BYTES), dtype=np.uint8)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index, period_ms=100):
        self.index, self.proc, self.lines, self.period_ms = index, None, [], period_ms

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", str(self.period_ms)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def mark(self):
        """Number of samples so far (to cut one timed region out of a longer sampling window)."""
        return len(self.lines)

    @staticmethod
    def summarise(lines):
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in lines:
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
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_mhz_min": min(sm) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "power_w_max": max(power) if power else None,
                "samples": len(sm), "reasons": sorted(reasons)}

    def stop(self):
        if not self.proc:
            return
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.proc = None


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


def simd_proxy_frames_per_s(threads, seconds=3.0):
    """What an optimised SIMD CPU library does with the same chain on this host: OpenCV's cvtColor(NV12 -> RGB) + bilinear
    resize + float normalise (AVX2 rows, one OpenCV thread per Python thread, the GIL released inside the calls). NOT the
    reference's arithmetic (OpenCV's fixed point differs in the last bits), so it is no parity arm -- it is the stand-in for the
    AVX2 rows of the real libyuv (the oracle port has its own AVX2 rows for the full-frame passes since late round 2; this stays as an
    independent check of its speed). Runs in a child interpreter (scripts/simd_proxy.py):
    importing cv2 into a process that already holds torch + a CUDA context crashed on the GPU box. Returns (frames/s, what)
    or (None, why)."""
    import subprocess
    script = os.path.join(os.path.dirname(os.path.abspath(__file__)), "scripts", "simd_proxy.py")
    try:
        r = subprocess.run([sys.executable, script, str(W), str(H), str(DW), str(DH), str(threads), str(seconds)],
                           capture_output=True, text=True, timeout=seconds * 4 + 120)
        if r.returncode != 0:
            return None, "scripts/simd_proxy.py exited %d (%s)" % (r.returncode, (r.stderr or "").strip().splitlines()[-1:] or "")
        d = json.loads(r.stdout.strip().splitlines()[-1])
        return d["frames_per_s"], d["what"]
    except Exception as e:  # noqa: BLE001
        return None, "simd proxy unavailable (%s)" % type(e).__name__


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


def _cpus_near_gpu(local):
    """(cpu id set, how) of the host CPUs closest to GPU `local`. Three sources, first that answers:
    NVML's ideal-CPU mask (what `nvidia-smi topo -m` prints as CPU Affinity), sysfs local_cpulist of the PCI function,
    the topo matrix itself."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        ncpu = os.cpu_count() or 64
        words = (ncpu + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        ids = {w * 64 + b for w, m in enumerate(mask) for b in range(64) if (int(m) >> b) & 1}
        node = None
        try:
            node = pynvml.nvmlDeviceGetNumaNodeId(h)
        except Exception:  # noqa: BLE001
            pass
        if ids:
            return ids, "nvml cpu affinity" + ("" if node is None else ", numa node %d" % node)
    except Exception:  # noqa: BLE001
        pass
    try:
        import torch
        bus = str(torch.cuda.get_device_properties(local).pci_bus_id).lower()
        if len(bus.split(":")[0]) == 8:
            bus = bus[4:]
        cpus = open("/sys/bus/pci/devices/" + bus + "/local_cpulist").read().strip()
        ids = set()
        for part in cpus.split(","):
            if "-" in part:
                a, b = part.split("-")
                ids.update(range(int(a), int(b) + 1))
            elif part:
                ids.add(int(part))
        if ids:
            return ids, "sysfs local_cpulist"
    except Exception:  # noqa: BLE001
        pass
    try:
        out = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout
        for ln in out.splitlines():
            if ln.startswith("GPU%d\t" % local) or ln.startswith("GPU%d " % local):
                for tok in ln.split():
                    if re.fullmatch(r"\d+(-\d+)?(,\d+(-\d+)?)*", tok) and ("-" in tok or "," in tok):
                        ids = set()
                        for part in tok.split(","):
                            a, _, b = part.partition("-")
                            ids.update(range(int(a), int(b or a) + 1))
                        return ids, "nvidia-smi topo -m"
    except Exception:  # noqa: BLE001
        pass
    return None, "no source answered (nvml, sysfs, nvidia-smi topo)"


def bind_near_gpu(local):
    """Pin this rank's threads to the CPUs of its GPU's NUMA node before any pinned allocation, so the frames and
    tensors of the e2e path sit in the memory closest to the GPU's PCIe root (one host thread per GPU).
    Returns a short description for the JSON line; never fatal."""
    if os.environ.get("BENCH_NO_AFFINITY") == "1":
        return "off"
    global _ALL_CPUS
    try:
        ids, how = _cpus_near_gpu(local)
        _ALL_CPUS = os.sched_getaffinity(0)
        if not ids:
            return "unavailable: " + how
        ids &= _ALL_CPUS
        if not ids:
            return "unavailable: affinity mask outside this process's cpus (%s)" % how
        if ids == _ALL_CPUS:
            return "single domain: all %d cpus are local to gpu %d (%s)" % (len(ids), local, how)
        os.sched_setaffinity(0, ids)
        return "bound to %d of %d cpus (%s)" % (len(ids), len(_ALL_CPUS), how)
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


class Events:
    """CUDA events through the shim (they time the processor's own launch stream, not torch's)."""

    def __init__(self):
        import ctypes as C
        from band_b200 import cuda_shim as S
        self.C, self.S, self.L = C, S, S.lib()

    def new(self):
        e = self.C.c_void_p()
        self.S.check(self.L.bandcu_event_create(self.C.byref(e)), "event_create")
        return e

    def record(self, e, stream):
        self.S.check(self.L.bandcu_event_record(e, stream), "event_record")

    def ms(self, a, b):
        out = self.C.c_float()
        self.S.check(self.L.bandcu_event_elapsed_ms(a, b, self.C.byref(out)), "elapsed")
        return float(out.value)


def run_ours(args):
    import ctypes as C
    import torch
    from band_b200 import band_c as B
    from band_b200 import hotpath
    from band_b200 import shard
    from band_b200 import traffic_model as T

    rank, local, world, dist = dist_setup(args.gpus)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product has no CPU fallback")
    torch.cuda.set_device(local)
    affinity = bind_near_gpu(local)
    hotpath.set_device(local)
    L = B.lib()
    EV = Events()
    ops = hotpath.chain_ops(DW, DH, True, MEAN, STD)
    F = args.frames
    K = args.steps
    Wm = max(args.warmup, 3)

    # ---- resident batch: F frames + F tensors in HBM (F * 3.1 MB >> 126 MB of L2: every step streams from HBM)
    rb = hotpath.ResidentBatch(ops, F, W, H, "NV12", np.float32, (1, DH, DW, 3), device=local)
    uniq = synth_frames(16, seed=0x42414E44 + rank)
    for i in range(F):
        if L.BandCopyToDevice(rb.dev_frames + i * rb.frame_pitch, uniq[i % 16].ctypes.data, FRAME_BYTES) != B.kBandOk:
            raise SystemExit("upload failed: " + B.last_error())
    rb.step()  # compiles the plan; creates the stream
    assert rb.is_fused, "north-star chain did not lower to the single fused launch"
    stream = rb.proc.stream
    for _ in range(Wm - 1):
        rb.step()

    evs = [(EV.new(), EV.new()) for _ in range(K)]
    e_begin, e_end = EV.new(), EV.new()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)  # first nvidia-smi sample lands before the timed region opens
    sync_all(dist, local)
    mark_timed = sampler.mark()
    launches_before = hotpath.launch_count()
    EV.record(e_begin, stream)
    for k in range(K):
        EV.record(evs[k][0], stream)
        st = rb.proc.process_batch_async(rb._ba, rb._ta, F)
        EV.record(evs[k][1], stream)
        if st != B.kBandOk:
            raise SystemExit("ProcessBatchAsync failed: " + B.last_error())
    EV.record(e_end, stream)
    rb.proc.synchronize()
    sync_all(dist, local)
    gpu_launches = hotpath.launch_count() - launches_before
    total_ms = barrier_max(dist, local, EV.ms(e_begin, e_end))
    kernel_ms = float(np.mean([EV.ms(a, b) for a, b in evs]))
    value = F * world * K / (total_ms * 1e-3)

    # ---- sustained leg: the same step back to back for >= args.sustain_seconds (not part of the K timed steps)
    n_sus = max(K, int(np.ceil(args.sustain_seconds * 1e3 / max(kernel_ms, 1e-3))))
    s0, s1 = EV.new(), EV.new()
    sync_all(dist, local)
    mark_sus = sampler.mark()
    EV.record(s0, stream)
    for _ in range(n_sus):
        if rb.proc.process_batch_async(rb._ba, rb._ta, F) != B.kBandOk:
            raise SystemExit("ProcessBatchAsync failed: " + B.last_error())
    EV.record(s1, stream)
    rb.proc.synchronize()
    sus_ms = barrier_max(dist, local, EV.ms(s0, s1))
    mark_sus_end = sampler.mark()
    sync_all(dist, local)
    if rank == 0:
        sampler.stop()
    clocks = ClockSampler.summarise(sampler.lines[mark_timed:]) if rank == 0 else None
    if clocks is not None:
        clocks["window"] = "the K timed steps plus the sustained leg that follows them (%.1f s); the timed steps alone last %.0f ms" % (
            sus_ms * 1e-3, total_ms)
    sus_clocks = ClockSampler.summarise(sampler.lines[mark_sus:mark_sus_end]) if rank == 0 else None
    sustained = {"value": F * world * n_sus / (sus_ms * 1e-3), "unit": "frames/s", "steps": n_sus, "seconds": sus_ms * 1e-3,
                 "ms_per_step": sus_ms / n_sus, "vs_timed_steps": (F * world * n_sus / (sus_ms * 1e-3)) / value,
                 "clocks": sus_clocks}

    # ---- parity: every rank checks one frame of ITS timed batch against the oracle (the checker, outside any timed region)
    mismatch = 0
    checked = 0
    if not args.no_check:
        from oracle import oracle as O
        o_ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, DW, DH), (O.OP_NORMALIZE, 0, 0, 0, 0, MEAN, STD)]
        n_chk = 2 if world == 1 else 1
        got_resident = rb.download(list(range(n_chk)))
        st, want = O.run_batch(o_ops, uniq[:n_chk], W, H, O.FMT_NV12, O.DT_F32, (1, DH, DW, 3))
        assert st == 0, st
        mismatch = int((got_resident.view(np.uint32) != want.view(np.uint32)).sum())
        checked = n_chk
    rb.close()

    # ---- e2e: host frames (pinned) -> BandImageProcessorProcessBatch -> host tensors (pinned), per step
    E = args.e2e_frames

    def e2e_leg(frame_base, steps, proc):
        """`steps` synchronous calls over E host frames starting at frame_base (pitch FRAME_BYTES) -> frames/s, launches."""
        tens = [B.Tensor(np.float32, (1, DH, DW, 3), device=False) for _ in range(E)]
        bufs = [B.Buffer().set_raw(frame_base + i * FRAME_BYTES, W, H, B.kBandNV12) for i in range(E)]
        ba = (C.c_void_p * E)(*[b.handle for b in bufs])
        ta = (C.c_void_p * E)(*[t.handle for t in tens])
        for _ in range(max(2, min(args.warmup, 3))):
            assert proc.process_batch_handles(ba, ta, E) == B.kBandOk, B.last_error()
        sync_all(dist, local)
        l0 = hotpath.launch_count()
        t0 = time.perf_counter()
        checksum = 0.0
        for _ in range(steps):
            assert proc.process_batch_handles(ba, ta, E) == B.kBandOk, B.last_error()
            # the call is synchronous: results are in host memory; touch one value per step like a consumer would
            checksum += float(np.ctypeslib.as_array(C.cast(tens[-1].data_ptr, C.POINTER(C.c_float)), shape=(4,))[0])
        ms = barrier_max(dist, local, (time.perf_counter() - t0) * 1e3)
        launches = hotpath.launch_count() - l0
        first = tens[0].numpy()
        for x in bufs + tens:
            x.close()
        return E * world * steps / (ms * 1e-3), launches, first

    host = L.BandHostAlloc(E * FRAME_BYTES)
    if not host:
        raise SystemExit("BandHostAlloc failed")
    hv = np.ctypeslib.as_array(C.cast(host, C.POINTER(C.c_uint8)), shape=(E, FRAME_BYTES))
    for i in range(E):
        hv[i] = uniq[i % 16]
    proc = B.ImageProcessor(ops)
    e2e_value, e2e_launches, got_e2e = e2e_leg(host, args.e2e_steps, proc)
    proc.close()
    L.BandHostFree(host)
    e2e_mismatch = None
    if not args.no_check:
        e2e_mismatch = int((got_e2e.view(np.uint32) != want[0].view(np.uint32)).sum())

    # ---- the unmodified caller: frames in plain malloc'ed (pageable) memory handed to BandBufferSetFromRawData
    e2e_pageable = None
    if world == 1 and not args.no_pageable:
        block = np.empty(E * FRAME_BYTES + 4096, dtype=np.uint8)
        off = (-block.ctypes.data) % 4096  # what a camera stack's allocator returns: page aligned, not locked
        pv = block[off:off + E * FRAME_BYTES].reshape(E, FRAME_BYTES)
        for i in range(E):
            pv[i] = uniq[i % 16]
        proc = B.ImageProcessor(ops)
        staged_value, staged_launches, got_p = e2e_leg(pv.ctypes.data, args.e2e_steps, proc)
        proc.close()
        B.host_registration_cache(True, 4 << 30)
        proc = B.ImageProcessor(ops)
        reg_value, reg_launches, got_r = e2e_leg(pv.ctypes.data, args.e2e_steps, proc)
        stats = B.host_registration_stats()
        proc.close()
        B.host_registration_cache(False)
        e2e_pageable = {
            "value": staged_value, "unit": "frames/s", "gpu_launches": staged_launches,
            "path": "pageable frames -> 32-frame chunks through plan-owned HBM staging (upload / compute / download streams)",
            "with_registration_cache": {"value": reg_value, "gpu_launches": reg_launches, "registrations": stats["registrations"],
                                        "registered_bytes": stats["registered_bytes"],
                                        "path": "BandHostRegistrationCacheConfigure(1): ranges page-locked on first sight "
                                                "(inside the warm-up calls), then read in place over PCIe like BandHostAlloc memory"},
            "frames_per_step": E, "steps": args.e2e_steps}
        if not args.no_check:
            e2e_pageable["mismatching_elements"] = int((got_p.view(np.uint32) != want[0].view(np.uint32)).sum()) + \
                int((got_r.view(np.uint32) != want[0].view(np.uint32)).sum())
        del block

    peak, peak_src = measured_peak()
    nec = T.necessary_yuv420_to_rgb(W, H, DW, DH, 12)
    NEC_BYTES = nec["necessary_bytes"]
    achieved_alg = ALG_BYTES * F / (kernel_ms * 1e-3) / 1e9
    achieved_nec = NEC_BYTES * F / (kernel_ms * 1e-3) / 1e9
    cap = load_ncu_capture()
    roofline = {"bound": "hbm", "achieved": achieved_nec, "peak": peak, "unit": "GB/s", "frac": achieved_nec / peak,
                "bytes_basis": "necessary: %d sampled luma rows + %d sampled chroma rows x 1920 B (every 32-byte sector of a sampled "
                               "row is touched: 8.6 px step) + the f32 tensor = %d B/frame (band_b200/traffic_model.py)"
                               % (nec["luma_rows"], nec["chroma_rows"], NEC_BYTES),
                "necessary_bytes_per_frame": NEC_BYTES, "bytes_per_launch": NEC_BYTES * F,
                "achieved_algorithmic": achieved_alg, "frac_algorithmic": achieved_alg / peak,
                "bytes_per_frame_algorithmic": ALG_BYTES, "bytes_per_launch_algorithmic": ALG_BYTES * F,
                "traffic": args.traffic if args.traffic is not None else (cap["bytes_per_frame"] * F if cap else None),
                "peak_source": peak_src, "kernel": KERNEL, "kernel_ms": kernel_ms, "frames_per_launch": F}
    if cap:
        roofline["traffic_source"] = {"file": cap["file"], "git_blob": cap["git_blob"], "frames_per_launch_in_capture": cap["frames_per_launch"],
                                      "launches_averaged": cap["launches"], "dram_bytes_per_frame": cap["bytes_per_frame"],
                                      "how": "ncu --set full: (dram__bytes_read.sum + dram__bytes_write.sum) / frames, scaled to this launch's F"}
        roofline["traffic_over_necessary"] = cap["bytes_per_frame"] / NEC_BYTES
        roofline["frac_on_ncu_traffic"] = cap["bytes_per_frame"] * F / (kernel_ms * 1e-3) / 1e9 / peak
        roofline["issue_active_pct"] = cap["issue_active_pct"]
        roofline["alu_pipe_pct"] = cap["alu_pipe_pct"]
    total_mismatch = int(shard.sum_over_ranks(dist, mismatch, device="cuda:%d" % local))
    total_checked = int(shard.sum_over_ranks(dist, checked, device="cuda:%d" % local))
    line = {
        "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": Wm,
        "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8 fixed-point (6/7/8-bit) -> f32 store", "data": "synthetic (seeded uniform bytes)",
        "config": {"workload": WORKLOAD, "frames_per_step_per_gpu": F, "bytes_per_frame_algorithmic": ALG_BYTES,
                   "bytes_per_frame_necessary": NEC_BYTES,
                   "l2_policy": "inputs larger than L2 (%.1f GB of frames per step per GPU vs 126 MB L2)" % (F * FRAME_BYTES / 1e9),
                   "parallelism": "frames sharded across %d GPU(s), no collective" % world,
                   "timing": "CUDA events on the processor's launch stream, max over ranks"},
        "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": E * FRAME_BYTES * world,
                "d2h_bytes_per_step": E * OUT_BYTES * world, "frames_per_step_per_gpu": E, "steps": args.e2e_steps,
                "timing": "host clock around synchronous BandImageProcessorProcessBatch calls (pinned host buffers and tensors), max over ranks",
                "gpu_launches": e2e_launches, "host_affinity": affinity,
                "bytes_crossing_pcie_per_frame": nec["luma_bytes"] + nec["chroma_bytes"] + OUT_BYTES,
                "note": "h2d/d2h_bytes_per_step count the full frames and tensors handed to the call; the fused kernel reads the "
                        "pinned frames in place over PCIe (only the source rows the resize samples cross the bus) and pushes "
                        "the tensors into pinned host memory with cp.async.bulk stores -- no staging copy either way"},
        "gpu_launches": gpu_launches, "roofline": roofline, "sustained": sustained, "clocks": clocks,
        "parity": None if args.no_check else {"frames_checked": total_checked, "ranks_checked": world, "mismatching_elements": total_mismatch,
                                              "e2e_mismatching_elements": e2e_mismatch,
                                              "oracle": "restated libyuv C paths; YUV->RGB arithmetic unpinned by any reference-held vector (DESIGN 5)"},
    }
    if e2e_pageable is not None:
        line["e2e_pageable"] = e2e_pageable
    unbind()
    # ---- secondary: the other BASELINE configs, by ONE process through the sharded API over the `world` GPUs of the box
    cpu_group = None
    if dist is not None:
        cpu_group = dist.new_group(backend="gloo")  # a host-side wait: no NCCL kernel spins on the GPUs rank 0 is measuring
        dist.barrier(group=cpu_group)
    if rank == 0 and not args.no_secondary:
        try:
            line["secondary"] = secondary_cases(world, peak, args)
        except Exception as e:  # noqa: BLE001 -- the headline must still print
            line["secondary"] = [{"case": "secondary leg failed", "error": "%s: %s" % (type(e).__name__, e)}]
        hotpath.set_device(local)
    if dist is not None:
        dist.barrier(group=cpu_group)

    if rank == 0 and world == 1 and not args.no_cpu:
        th = host_threads()
        fps1, _ = cpu_frames_per_s(4, th)                     # short probe to size the sample
        per_thread = max(4, int(fps1 / th * args.cpu_seconds))
        fps, secs = cpu_frames_per_s(per_thread, th)
        line["cpu_baseline"] = {"value": fps, "unit": "frames/s", "cores": th, "kind": "port",
                                "sample": "%d frames of the same workload (%d per thread x %d threads), %.1f s; restated C with AVX2 rows for the full-frame passes (4:2:0 -> RGB, RGB24 -> ARGB, InterpolateRow), bit-identical to its scalar form"
                                          % (per_thread * th, per_thread, th, secs)}
        proxy, what = simd_proxy_frames_per_s(th)
        line["cpu_baseline"]["simd_proxy"] = {
            "value": proxy, "unit": "frames/s", "what": what,
            "note": "an independent optimised CPU library on the same threads, as a sanity check of the port's speed (the real "
                    "libyuv is absent here); not bit-comparable, no parity arm"}
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        emit(line)


def _time_sharded(batches, reps):
    """Host clock around `reps` rounds of: enqueue every batch on every shard, then join all shards. Whole-job seconds
    per round (includes the host-thread hand-off a caller of the sharded API pays)."""
    for _ in range(2):
        for b in batches:
            b.step_async()
        for b in batches:
            b.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        for b in batches:
            b.step_async()
    for b in batches:
        b.synchronize()
    return (time.perf_counter() - t0) / reps


def secondary_cases(world, peak, args):
    """BASELINE configs C1..C5 through BandImageProcessorProcessBatchSharded over devices 0..world-1: frames
    and tensors of each shard's contiguous slice live on that shard's GPU. frames/s are whole-job; `frac` is per GPU on
    NECESSARY bytes (band_b200/traffic_model.py) against the measured HBM peak, `frac_algorithmic` on SURVEY 8d's figure."""
    import ctypes as C
    from band_b200 import band_c as B
    from band_b200 import hotpath
    from band_b200 import traffic_model as T
    devices = list(range(world))
    out = []
    rng = np.random.default_rng(0xC4)
    reps = 12

    def yuv_case(name, Wd, Hd, fmt, dw, dh, f32, per_gpu, kernel, pre_ops=(), region=None, swap=False):
        fb = B.frame_bytes(Wd, Hd, B.FORMATS[fmt])
        uniq = rng.integers(0, 256, size=(4, fb), dtype=np.uint8)
        ops = list(pre_ops) + hotpath.chain_ops(dw, dh, f32)
        sb = hotpath.ShardedResidentBatch(ops, per_gpu * world, Wd, Hd, fmt, np.float32 if f32 else np.uint8, (1, dh, dw, 3),
                                          devices, uniq)
        sec = _time_sharded([sb], reps)
        fps = sb.n / sec
        obpp = 12 if f32 else 3
        nec = T.necessary_yuv420_to_rgb(Wd, Hd, dw, dh, obpp, interleaved=fmt in ("NV12", "NV21"), region=region, swap_axes=swap)
        alg = fb + dw * dh * obpp
        sb.close()
        return {"case": name, "value": fps, "unit": "frames/s", "frames_per_gpu": per_gpu, "kernel": kernel,
                "necessary_bytes_per_frame": nec["necessary_bytes"], "frac": fps / world * nec["necessary_bytes"] / 1e9 / peak,
                "frac_algorithmic": fps / world * alg / 1e9 / peak, "timing": "host clock around the sharded call, %d rounds" % reps}

    if world == 1:
        out += c1_latency_cases()
        out.append(yuv_case("C1 batch: 640x480 NV21 -> 224x224 f32", 640, 480, "NV21", 224, 224, True, 4096,
                            "fast_nv_bilinear_down_kernel<float, kChromaVU>"))
    out.append(yuv_case("C2: 1080p NV12 -> 300x300 u8", 1920, 1080, "NV12", 300, 300, False, 1024,
                        "fast_nv_bilinear_down_kernel<uint8_t, kChromaUV>"))
    out.append(yuv_case("C3: 4K YV12 -> Rotate 90 -> Crop 2160x2160 -> 512x512 f32", 3840, 2160, "YV12", 512, 512, True, 128,
                        "tile_yuv_bilinear_down_kernel<float, planar chroma, quarter-turn>",
                        pre_ops=[(B.BAND_ROTATE, 90), (B.BAND_CROP, 0, 840, 2159, 2999)], region=(840, 0, 2160, 2160), swap=True))

    # C4: frame i of the batch goes to target i % 3; each target's member list is cut into contiguous slices, one per GPU,
    # so every GPU gets the same 224 / 300 / 513 mix. Three sharded processors, queued together, one join.
    targets = [(224, True), (300, False), (513, True)]
    for resident, per_gpu in ((True, 512), (False, 96)):
        n_total = per_gpu * world
        fb = B.frame_bytes(1920, 1080, B.kBandNV21)
        uniq = rng.integers(0, 256, size=(4, fb), dtype=np.uint8)
        batches, nec_sum, alg_sum = [], 0, 0
        for t, (d, f32) in enumerate(targets):
            members = len(range(t, n_total, 3))
            sb = hotpath.ShardedResidentBatch(hotpath.chain_ops(d, d, f32), members, 1920, 1080, "NV21",
                                              np.float32 if f32 else np.uint8, (1, d, d, 3), devices, uniq, resident=resident)
            batches.append(sb)
            obpp = 12 if f32 else 3
            nec_sum += members * T.necessary_yuv420_to_rgb(1920, 1080, d, d, obpp)["necessary_bytes"]
            alg_sum += members * (fb + d * d * obpp)
        sec = _time_sharded(batches, reps if resident else 6)
        fps = n_total / sec
        for sb in batches:
            sb.close()
        line = {"case": "C4 mixed: %d frames 1080p NV21 -> {224 f32, 300 u8, 513 f32}[i %% 3], %d per GPU, %s"
                        % (n_total, per_gpu, "resident in HBM" if resident else "e2e from pinned host frames into pinned host tensors"),
                "value": fps, "unit": "frames/s", "frames_per_gpu": per_gpu, "kernel": "fast_nv_bilinear_down_kernel x3 (one launch per target and GPU)",
                "timing": "host clock around three ProcessBatchShardedAsync calls + three joins"}
        if resident:
            line["frac"] = nec_sum / sec / world / 1e9 / peak
            line["frac_algorithmic"] = alg_sum / sec / world / 1e9 / peak
        out.append(line)

    # C5 extremes: RGBA S^2 -> RGB D^2 through the default AutoConvert processor (empty builder)
    for S_, D_, kernel in ((2048, 64, "fused_generic_kernel (ratio 32: sparse gathers)"),
                           (2048, 1024, "packed_bilinear_kernel (exact 2x box)"),
                           (64, 2048, "packed_up_rows_x16_kernel")):
        per_gpu = 256 if D_ == 64 else 64   # the 2048 -> 64 launch is ~12 us per 64 frames: give the hand-off something to amortise
        uniq = rng.integers(0, 256, size=(2, 4 * S_ * S_), dtype=np.uint8)
        sb = hotpath.ShardedResidentBatch([], per_gpu * world, S_, S_, "RGBA", np.uint8, (1, D_, D_, 3), devices, uniq)
        sec = _time_sharded([sb], reps)
        fps = sb.n / sec
        sb.close()
        nec = T.necessary_packed(S_, D_, 4, 3)["necessary_bytes"]
        alg = 4 * S_ * S_ + 3 * D_ * D_
        out.append({"case": "C5: RGBA %dx%d -> RGB %dx%d u8 (AutoConvert)" % (S_, S_, D_, D_), "value": fps, "unit": "frames/s",
                    "frames_per_gpu": per_gpu, "kernel": kernel, "necessary_bytes_per_frame": nec,
                    "frac": fps / world * nec / 1e9 / peak, "frac_algorithmic": fps / world * alg / 1e9 / peak,
                    "timing": "host clock around the sharded call, %d rounds" % reps})
    return out


def c1_latency_cases():
    """BASELINE configs[0]: ONE 640x480 NV21 frame -> 224x224 f32, latency of a synchronous BandImageProcessorProcess and of
    the same call recorded into a CUDA graph (frame and tensor resident in HBM)."""
    from band_b200 import band_c as B
    from band_b200 import hotpath
    Wd, Hd = 640, 480
    proc = B.ImageProcessor(hotpath.chain_ops(224, 224, True))
    t = B.Tensor(np.float32, (1, 224, 224, 3), device=True)
    fb = B.frame_bytes(Wd, Hd, B.kBandNV21)
    frame = np.random.default_rng(5).integers(0, 256, size=fb, dtype=np.uint8)
    L = B.lib()
    dev = L.BandDeviceAlloc(fb)
    L.BandCopyToDevice(dev, frame.ctypes.data, fb)
    buf = B.Buffer().set_raw(dev, Wd, Hd, B.kBandNV21, device=True)
    out = []

    def lat(call, what):
        for _ in range(30):
            assert call() == B.kBandOk, B.last_error()
        ts = []
        for _ in range(400):
            t0 = time.perf_counter()
            call()
            ts.append((time.perf_counter() - t0) * 1e6)
        ts = np.array(ts)
        out.append({"case": "C1 latency: one 640x480 NV21 frame -> 224x224 f32, " + what, "value": float(np.median(ts)), "unit": "us",
                    "p10": float(np.percentile(ts, 10)), "p90": float(np.percentile(ts, 90)), "higher_is_better": False,
                    "kernel": "fast_nv_bilinear_down_kernel<float, kChromaVU>", "timing": "host clock around the synchronous call, median of 400"})

    lat(lambda: proc.process(buf, t), "BandImageProcessorProcess")
    rec = proc.record_batch([buf], [t])
    lat(lambda: proc.replay(rec), "CUDA-graph replay (BandRecordedBatchReplay)")
    proc.release(rec)
    buf.close(); t.close(); proc.close()
    L.BandDeviceFree(dev)
    return out


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
            "cpu_baseline": {"value": value, "unit": "frames/s", "cores": th, "kind": "port",
                             "sample": sample + "; restated C with AVX2 rows for the full-frame passes (4:2:0 -> RGB, RGB24 -> ARGB, "
                                                "InterpolateRow), bit-identical to its scalar form (tests/test_oracle_golden.py)"},
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
    ap.add_argument("--e2e-steps", type=int, default=24)
    ap.add_argument("--sustain-seconds", type=float, default=2.0)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--ref-step-seconds", type=float, default=4.0)
    ap.add_argument("--traffic", type=float, default=None, help="ncu dram bytes per launch, overriding the committed capture")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-check", action="store_true")
    ap.add_argument("--no-secondary", action="store_true")
    ap.add_argument("--no-pageable", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
