#!/usr/bin/env python
"""Session fuzzer: ONE processor called many times the way a camera loop does -- the same handles refilled in place, new
handles over the same memory, another 4:2:0 layout of the same size, another kind of memory (pageable, pinned, HBM,
padded strides), another batch size, host or HBM tensors -- with a persistent oracle processor fed the same sequence.
Every call's output must equal the oracle's bit for bit and the accept / reject verdict must agree.

scripts/fuzz_chains.py builds a fresh processor per case; the host layer's state (plan cache, repeat-call replay,
resolved-pointer tables, staging buffers, the registration cache) only shows its bugs across calls.

  python scripts/fuzz_sessions.py --sessions 300 --seed 1      # on a B200
  python scripts/fuzz_sessions.py --only 17                    # replay one session verbosely
"""
import argparse
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "scripts"))

from oracle import oracle as O         # noqa: E402  (test infrastructure: the checker)
import fuzz_chains as F                # noqa: E402

MEMS = ["host", "resident", "pinned", "padded_host", "padded_resident"]


class OracleSession:
    """A persistent oracle processor: operators keep their auto-created intermediates between calls like the
    reference's do (image_operator.cc CreateOutput runs once)."""

    def __init__(self, ops):
        self.L = O.lib()
        st = C.c_int(0)
        self.descs = O.make_descs(ops)
        self.proc = self.L.orc_processor_build(self.descs, len(ops), C.byref(st))
        assert self.proc, st.value

    def run(self, frames, W, H, fmt, out_dt, dims):
        outs = []
        cdims = (C.c_int * len(dims))(*dims)
        for f in frames:
            out = np.zeros(dims, dtype=O._NP_DT[out_dt])
            ib = self.L.orc_buffer_from_raw(f.ctypes.data, W, H, fmt, O.DT_U8)
            ob = self.L.orc_buffer_from_tensor(out.ctypes.data, out_dt, cdims, len(dims))
            try:
                st = self.L.orc_processor_process(self.proc, ib, ob)
            finally:
                self.L.orc_buffer_free(ib)
                self.L.orc_buffer_free(ob)
            if st != 0:
                return st, None
            outs.append(out)
        return 0, np.stack(outs)

    def close(self):
        self.L.orc_processor_free(self.proc)


class FrameSet:
    """n frames of one layout in one kind of memory, refillable in place, with their Buffer handles."""

    def __init__(self, B, W, H, fmt, mem, n, pads, misalign):
        self.B, self.L = B, B.lib()
        self.W, self.H, self.fmt, self.mem, self.n, self.pads = W, H, fmt, mem, n, pads
        self.fb = O.frame_bytes(W, H, fmt)
        self.host = self.dev = None
        self.arrays = []
        self.bufs = []
        self.device = mem.endswith("resident")
        if mem in ("host", "resident", "pinned"):
            self.pitch = (self.fb + 255) // 256 * 256
            self.mis = misalign if mem == "pinned" else 0
            if mem == "pinned":
                self.host = self.L.BandHostAlloc(self.pitch * n + 64)
            elif mem == "resident":
                self.dev = self.L.BandDeviceAlloc(self.pitch * n)
            else:
                self.arrays = [np.zeros(self.fb, dtype=np.uint8) for _ in range(n)]
        else:
            lay = F._padded_layout(np.zeros(self.fb, dtype=np.uint8), W, H, fmt, pads[0], pads[1])
            self.size = lay[0].size
            self.pitch = (self.size + 255) // 256 * 256
            if self.device:
                self.dev = self.L.BandDeviceAlloc(self.pitch * n)
            else:
                self.arrays = [np.zeros(self.size, dtype=np.uint8) for _ in range(n)]
        self.make_handles()

    def base(self, i):
        if self.mem == "pinned":
            return self.host + self.mis + i * self.pitch
        if self.device:
            return self.dev + i * self.pitch
        return self.arrays[i].ctypes.data

    def make_handles(self):
        for b in self.bufs:
            b.close()
        self.bufs = []
        B = self.B
        for i in range(self.n):
            if self.mem in ("host", "resident", "pinned"):
                self.bufs.append(B.Buffer().set_raw(self.base(i), self.W, self.H, self.fmt, device=self.device))
            else:
                _, yo, uo, vo, ys, cs, ps = F._padded_layout(np.zeros(self.fb, dtype=np.uint8), self.W, self.H, self.fmt, *self.pads)
                b = self.base(i)
                self.bufs.append(B.Buffer().set_yuv(b + yo, b + uo, b + vo, self.W, self.H, ys, cs, ps, self.fmt, device=self.device))

    def fill(self, frames):
        """New pixels into the SAME memory (what a camera does with its buffers)."""
        for i in range(self.n):
            if self.mem in ("host", "resident", "pinned"):
                src, size = frames[i], self.fb
            else:
                src, size = F._padded_layout(frames[i], self.W, self.H, self.fmt, *self.pads)[0], self.size
            if self.device:
                assert self.L.BandCopyToDevice(self.base(i), src.ctypes.data, size) == self.B.kBandOk
            elif self.mem == "pinned":
                C.memmove(self.base(i), src.ctypes.data, size)
            else:
                self.arrays[i][:] = src

    def close(self):
        for b in self.bufs:
            b.close()
        if self.host:
            self.L.BandHostFree(self.host)
        if self.dev:
            self.L.BandDeviceFree(self.dev)


def family(fmt):
    return [O.FMT_NV12, O.FMT_NV21] if fmt in (O.FMT_NV12, O.FMT_NV21) else [O.FMT_YV12, O.FMT_YV21]


def run_session(seed, verbose=False):
    """-> list of failure strings (empty = the whole session agreed)."""
    import gpu_util as G
    from band_b200 import band_c as B
    rng = np.random.default_rng(seed * 31 + 5)
    c = F.make_staged_case(seed * 3)
    if c["W"] * c["H"] > 1300 * 800:   # sessions make many calls: keep frames moderate
        c["W"], c["H"] = 640, 360
        c = dict(c, ops=[o for o in c["ops"] if o[0] != O.OP_CROP])
    shape = F.arith_shape(c)
    if shape is None or shape[0] < 2 or shape[1] < 2:
        return []
    dims = (1, shape[1], shape[0], 3)
    W, H = c["W"], c["H"]
    ops = c["ops"]
    fails = []
    osess = OracleSession(ops)
    proc = B.ImageProcessor(G.to_capi_ops(ops))
    fmt, mem, n = c["fmt"], c["mem"], c["n"]
    pads, mis = c["pads"], c["misalign"]
    fs = FrameSet(B, W, H, fmt, mem, n, pads, mis)
    tens_device = bool(rng.random() < 0.5)
    tens = [B.Tensor(G.NP_OF_ORC[c["out_dt"]], dims, device=tens_device) for _ in range(n)]
    if verbose:
        print(F.describe(c), "dims", dims)
    try:
        steps = int(rng.integers(5, 11))
        for k in range(steps):
            act = str(rng.choice(["refill", "refill", "refill", "new_handles", "layout", "memory", "batch", "tensors", "fmt_family"])) if k else "first"
            # A processor whose FIRST operator is Crop / Rotate / Flip keeps an intermediate in the first frame's layout. The
            # reference then accepts any other 4:2:0 layout (buffer.cc:367-388) but only handles the planar pair correctly
            # (planes resolved by name); semi-planar pairs get U and V exchanged or written one byte off, planar <-> semi-planar
            # overruns. The product follows the planar pair and refuses the rest (image_operator.cc LayoutsAgree): not compared.
            index_first = ops[0][0] in (O.OP_CROP, O.OP_ROTATE, O.OP_FLIP)
            if index_first and (act == "fmt_family" or (act == "layout" and fmt in (O.FMT_NV12, O.FMT_NV21))):
                act = "refill"
            if act == "new_handles":
                fs.make_handles()
            elif act == "layout":      # U/V order swapped: same size, same family
                fmt = [g for g in family(fmt) if g != fmt][0]
            elif act == "fmt_family":  # planar <-> semi-planar of the same size
                fmt = int(rng.choice([g for g in F.YUV if g not in family(fmt)]))
            elif act == "memory":
                mem = str(rng.choice(MEMS))
                pads, mis = (int(rng.choice([0, 16, 64, 8, 3])), int(rng.choice([0, 16, 32, 6, 1]))), int(rng.choice([0, 0, 4, 1]))
            elif act == "batch":
                n = int(rng.choice([1, 2, 3, 5, 8]))
            elif act == "tensors":
                tens_device = not tens_device
            if act in ("layout", "fmt_family", "memory", "batch"):
                fs.close()
                fs = FrameSet(B, W, H, fmt, mem, n, pads, mis)
            if act in ("batch", "tensors"):
                for t in tens:
                    t.close()
                tens = [B.Tensor(G.NP_OF_ORC[c["out_dt"]], dims, device=tens_device) for _ in range(n)]
            frames = G.rand_frames(O, n, W, H, fmt, seed=seed * 1000 + k)
            fs.fill(frames)
            st_o, want = osess.run(frames, W, H, fmt, c["out_dt"], dims)
            st_g = proc.process_batch(fs.bufs, tens) if n != 1 else proc.process(fs.bufs[0], tens[0])
            what = "session %d step %d (%s): %s %dx%d x%d %s, %s tensors" % (seed, k, act, F.FMT_NAMES[fmt], W, H, n, mem,
                                                                              "HBM" if tens_device else "host")
            if verbose:
                print("  ", what, "oracle", st_o, "product", st_g)
            if (st_o == 0) != (st_g == B.kBandOk):
                fails.append("%s: oracle status %d, product status %d (%s)" % (what, st_o, st_g, B.last_error()))
                break  # the two processors' states have diverged
            if st_o != 0:
                continue
            got = np.stack([t.numpy() for t in tens])
            bad = int((np.ascontiguousarray(got).view(np.uint8) != np.ascontiguousarray(want).view(np.uint8)).sum())
            if bad:
                fails.append("%s: %d of %d bytes differ" % (what, bad, got.nbytes))
    finally:
        fs.close()
        for t in tens:
            t.close()
        proc.close()
        osess.close()
    if fails:
        fails.insert(0, "chain: " + F.describe(c))
    return fails


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sessions", type=int, default=200)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--only", type=int, default=None)
    ap.add_argument("--max-fail", type=int, default=20)
    a = ap.parse_args()
    if a.only is not None:
        for line in run_session(a.only, verbose=True):
            print("FAIL", line)
        return 0
    bad = []
    t0 = time.time()
    for s in range(a.seed, a.seed + a.sessions):
        try:
            fails = run_session(s)
        except Exception as e:  # noqa: BLE001
            fails = ["session %d: exception %s: %s" % (s, type(e).__name__, e)]
        if fails:
            bad.append(s)
            for line in fails:
                print("FAIL", line, flush=True)
            if len(bad) >= a.max_fail:
                break
    print("fuzz_sessions: sessions %d..%d in %.0f s; failing sessions: %s" % (a.seed, a.seed + a.sessions - 1, time.time() - t0, bad))
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
