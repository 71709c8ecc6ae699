#!/usr/bin/env python
"""Differential fuzzer: random operator chains on random frames through the product's C API (CUDA) and through the CPU
oracle; every byte of every output (or the ok / not-ok verdict when the reference rejects the chain) must agree.

The hand-written parity tests pin the cases somebody thought of; this walks the rest of the space: every source format,
odd and tiny dimensions, every order of Crop / Rotate / Flip / ColorSpaceConvert / Resize the builder accepts, exact-ratio
and free resizes in both directions, Normalize / DataTypeConvert into every tensor type, single frames and batches,
host and HBM-resident buffers. A failing case prints the one-line recipe that replays it (--only SEED).

  python scripts/fuzz_chains.py --cases 3000 --seed 1        # on a B200
  python scripts/fuzz_chains.py --only 1234                  # replay one case verbosely
"""
import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import oracle as O         # noqa: E402  (test infrastructure: the checker)

FMT_NAMES = {O.FMT_GRAY: "GRAY", O.FMT_RGB: "RGB", O.FMT_RGBA: "RGBA", O.FMT_YV12: "YV12", O.FMT_YV21: "YV21",
             O.FMT_NV21: "NV21", O.FMT_NV12: "NV12"}
YUV = (O.FMT_YV12, O.FMT_YV21, O.FMT_NV21, O.FMT_NV12)
CHANNELS = {O.FMT_GRAY: 1, O.FMT_RGB: 3, O.FMT_RGBA: 4}
OP_NAMES = {O.OP_CROP: "Crop", O.OP_RESIZE: "Resize", O.OP_ROTATE: "Rotate", O.OP_FLIP: "Flip", O.OP_CSC: "CSC",
            O.OP_NORMALIZE: "Normalize", O.OP_DTC: "DTC", O.OP_AUTO: "Auto"}


def pick_dims(rng):
    r = rng.random()
    if r < 0.25:
        return int(rng.integers(1, 24)), int(rng.integers(1, 24))
    if r < 0.75:
        return int(rng.integers(8, 200)), int(rng.integers(8, 160))
    if r < 0.93:
        w, h = int(rng.integers(100, 700)), int(rng.integers(60, 500))
        if rng.random() < 0.5:
            w, h = w // 16 * 16 + 16, h // 2 * 2 + 2   # the aligned shapes the staged kernels take
        return w, h
    return int(rng.choice([640, 1280, 1920])), int(rng.choice([360, 480, 720]))


def pick_resize(rng, w, h):
    r = rng.random()
    if r < 0.30:   # exact ratios: libyuv's special cases (Down2, Down4, 3/4, 3/8, Up2, same size)
        num, den = [(1, 2), (1, 4), (3, 4), (3, 8), (2, 1), (1, 1), (1, 3), (4, 1)][int(rng.integers(0, 8))]
        return max(1, w * num // den), max(1, h * num // den)
    if r < 0.60:   # free down-scale
        return max(1, int(w * rng.uniform(0.05, 1.0))), max(1, int(h * rng.uniform(0.05, 1.0)))
    if r < 0.80:   # free up-scale (bounded output)
        f = rng.uniform(1.0, 4.0)
        return min(1024, max(1, int(w * f))), min(1024, max(1, int(h * rng.uniform(1.0, 4.0))))
    if r < 0.9:    # mixed
        return min(1024, max(1, int(w * rng.uniform(0.3, 2.0)))), min(1024, max(1, int(h * rng.uniform(0.3, 2.0))))
    return int(rng.choice([224, 300, 64, 128, 96])), int(rng.choice([224, 300, 64, 128, 96]))


VALID_TARGETS = {  # ColorSpaceConvert::ValidateOutput, image_operator.cc:367-399
    O.FMT_GRAY: [],
    O.FMT_RGB: [O.FMT_GRAY, O.FMT_NV21, O.FMT_NV12, O.FMT_YV12, O.FMT_YV21],
    O.FMT_RGBA: [O.FMT_RGB, O.FMT_RGB, O.FMT_GRAY, O.FMT_NV21, O.FMT_NV12, O.FMT_YV12, O.FMT_YV21],
}
for _f in YUV:
    VALID_TARGETS[_f] = [O.FMT_RGB, O.FMT_RGB, O.FMT_RGB, O.FMT_RGBA, O.FMT_GRAY] + [g for g in YUV if g != _f]


def make_staged_case(seed):
    """The shapes the staged kernels are built for -- 4:2:0 source, (Crop / Rotate / Flip) -> ColorSpaceConvert(RGB) ->
    Resize down -> (Normalize) in the orders that fold into one launch -- with the geometry left free: source sizes around the
    16- and 32-byte granules, even and odd crop boxes, output sizes from 8 to 640 (two passes above 512 columns), every
    layout, batches, pageable / pinned / HBM-resident / padded-stride frames."""
    rng = np.random.default_rng(seed)
    fmt = int(rng.choice(YUV))
    r = rng.random()
    if r < 0.45:
        W, H = int(rng.integers(4, 60)) * 16, int(rng.integers(20, 300)) * 2
    elif r < 0.75:
        W, H = int(rng.integers(20, 500)) * 2, int(rng.integers(20, 300)) * 2
    elif r < 0.85:
        W, H = int(rng.integers(33, 700)), int(rng.integers(33, 500))
    else:
        W, H = [(640, 480), (1280, 720), (1920, 1080), (1920, 1088), (3840, 2160)][int(rng.integers(0, 5))]
    pre = []
    w, h = W, H
    kinds = [O.OP_CROP, O.OP_ROTATE, O.OP_FLIP]
    rng.shuffle(kinds)
    for k in kinds[:int(rng.choice([0, 0, 1, 1, 2, 3]))]:
        if k == O.OP_CROP:
            x0, y0 = int(rng.integers(0, w // 2)), int(rng.integers(0, h // 2))
            x1, y1 = int(rng.integers(x0 + w // 4, w)), int(rng.integers(y0 + h // 4, h))
            if rng.random() < 0.8:
                x0, y0 = x0 // 2 * 2, y0 // 2 * 2
                x1, y1 = min(w - 1, x0 + max(2, (x1 - x0 + 1) // 2 * 2) - 1), min(h - 1, y0 + max(2, (y1 - y0 + 1) // 2 * 2) - 1)
            pre.append((k, x0, y0, x1, y1))
            w, h = x1 - x0 + 1, y1 - y0 + 1
        elif k == O.OP_ROTATE:
            a = int(rng.choice([90, 180, 270]))
            pre.append((k, a))
            if a != 180:
                w, h = h, w
        else:
            hv = int(rng.integers(1, 4))
            pre.append((k, hv & 1, hv >> 1))
    r = rng.random()
    if r < 0.7:
        dw, dh = max(2, int(w * rng.uniform(0.08, 1.0))), max(2, int(h * rng.uniform(0.08, 1.0)))
    elif r < 0.9:
        dw = dh = int(rng.choice([64, 96, 128, 224, 256, 299, 300, 320, 384, 512, 513, 640]))
    else:
        dw, dh = max(2, w // int(rng.choice([1, 2, 3, 4]))), max(2, h // int(rng.choice([1, 2, 3, 4])))
    dw, dh = min(dw, 1024), min(dh, 1024)
    post = []
    if rng.random() < 0.15:  # an index operator AFTER the conversion (does not fold into the row-staged kernel)
        post = [(O.OP_FLIP, 1, 0)] if not any(o[0] == O.OP_FLIP for o in pre) else []
    order = rng.random()
    if order < 0.8:
        ops = pre + [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, dw, dh)] + post
    else:  # conversion first: the index operators act on the RGB image
        ops = [(O.OP_CSC, O.FMT_RGB)] + pre + [(O.OP_RESIZE, dw, dh)] + post
    out_dt, tail = O.DT_U8, None
    r = rng.random()
    if r < 0.6:
        out_dt = O.DT_F32
        mean, std = [(127.5, 127.5), (0.0, 255.0), (128.0, 1.0), (103.5, 57.375)][int(rng.integers(0, 4))]
        tail = (O.OP_NORMALIZE, 0, 0, 0, 0, mean, std)
    elif r < 0.7:
        out_dt = O.DT_I8
        tail = (O.OP_NORMALIZE, 0, 0, 0, 0, 128.0, 1.0)
    elif r < 0.75:
        out_dt = O.DT_F32
        tail = (O.OP_DTC,)
    if tail:
        ops.append(tail)
    big = W * H > 1000 * 1000
    n = int(rng.choice([1, 2, 3, 7, 16])) if not big else int(rng.choice([1, 2, 3]))
    mem = str(rng.choice(["host", "resident", "pinned", "padded_host", "padded_resident", "resident", "pinned"]))
    return dict(seed=seed, fmt=fmt, W=W, H=H, ops=ops, mode="tensor", out_dt=out_dt, n=n, resident=(mem == "resident"),
                tail=tail, mem=mem, pads=(int(rng.choice([0, 16, 64, 8, 3])), int(rng.choice([0, 16, 32, 6, 1]))),
                misalign=int(rng.choice([0, 0, 0, 4, 1])))


def make_case(seed):
    if seed % 3 == 0:
        return make_staged_case(seed)
    return make_free_case(seed)


def make_free_case(seed):
    """A chain that is USUALLY valid (the shape and format are followed down the chain), with a few per cent of each kind of
    mistake the reference rejects: bad angles, crop boxes outside the image, conversions it does not have, a tensor of the
    wrong type."""
    rng = np.random.default_rng(seed)
    fmt = int(rng.choice(list(FMT_NAMES)))
    W, H = pick_dims(rng)
    if fmt in YUV and rng.random() < 0.8:
        W, H = max(2, W // 2 * 2), max(2, H // 2 * 2)
    mode = "tensor" if rng.random() < 0.7 else "image"
    kinds = [O.OP_CROP, O.OP_ROTATE, O.OP_FLIP, O.OP_CSC, O.OP_RESIZE]
    rng.shuffle(kinds)
    kinds = kinds[:int(rng.integers(0, 6))]
    if mode == "tensor" and fmt != O.FMT_RGB and O.OP_CSC not in kinds and fmt != O.FMT_GRAY and rng.random() < 0.92:
        kinds.insert(int(rng.integers(0, len(kinds) + 1)), O.OP_CSC)   # a tensor is an RGB image (buffer.cc:106-152)
    ops = []
    w, h, f = W, H, fmt
    for k in kinds:
        if k == O.OP_CROP:
            if rng.random() < 0.04 or w < 1 or h < 1:
                op = (k, int(rng.integers(0, max(1, w) + 3)), int(rng.integers(0, max(1, h) + 3)),
                      int(rng.integers(0, max(1, w) + 3)), int(rng.integers(0, max(1, h) + 3)))
            else:
                x0, y0 = int(rng.integers(0, w)), int(rng.integers(0, h))
                x1, y1 = int(rng.integers(x0, w)), int(rng.integers(y0, h))
                if rng.random() < 0.6:  # even origin and extent: what a YUV crop wants
                    x0, y0 = x0 // 2 * 2, y0 // 2 * 2
                    x1 = min(w - 1, x0 + max(2, (x1 - x0 + 1) // 2 * 2) - 1)
                    y1 = min(h - 1, y0 + max(2, (y1 - y0 + 1) // 2 * 2) - 1)
                op = (k, x0, y0, x1, y1)
            w, h = op[3] - op[1] + 1, op[4] - op[2] + 1
        elif k == O.OP_ROTATE:
            op = (k, int(rng.choice([90, 180, 270])) if rng.random() < 0.96 else int(rng.choice([0, 45, 360])))
            if op[1] in (90, 270):
                w, h = h, w
        elif k == O.OP_FLIP:
            hv = int(rng.integers(1, 4))
            op = (k, hv & 1, hv >> 1)
        elif k == O.OP_CSC:
            targets = VALID_TARGETS[f]
            if mode == "tensor" and O.FMT_RGB in targets and rng.random() < 0.85:
                t = O.FMT_RGB
            elif targets and rng.random() < 0.95:
                t = int(rng.choice(targets))
            else:
                t = int(rng.choice(list(FMT_NAMES)))
            op = (k, t)
            f = t
        else:
            op = (k,) + pick_resize(rng, max(1, w), max(1, h))
            w, h = op[1], op[2]
        ops.append(op)
    if mode == "tensor" and ops and ops[-1][0] == O.OP_FLIP and ops[-1][1] and ops[-1][2]:
        # a two-axis Flip writing the caller's buffer dereferences an intermediate that only CreateOutput makes
        # (image_operator.cc:299-309): the reference crashes, nothing to compare with
        ops[-1] = (O.OP_FLIP, 1, 0)
    tail = None
    out_dt = O.DT_U8
    if mode == "tensor":
        r = rng.random()
        if r < 0.5:
            # the reference casts (x - mean) to the output type BEFORE dividing (common_operator.cc:24-40): a difference the
            # type cannot hold is undefined behaviour in C++, so each type only gets means that keep it in range
            out_dt = int(rng.choice([O.DT_F32, O.DT_F32, O.DT_F32, O.DT_I8, O.DT_U8, O.DT_I16, O.DT_I32]))
            if out_dt == O.DT_U8:
                choices = [(0.0, 255.0), (0.0, 1.0), (0.0, 2.0)]
            elif out_dt == O.DT_I8:
                choices = [(128.0, 1.0), (127.5, 127.5), (127.5, 1.0), (128.0, 2.0)]
            else:
                choices = [(127.5, 127.5), (0.0, 255.0), (128.0, 1.0), (103.5, 57.375), (0.0, 1.0)]
            mean, std = choices[int(rng.integers(0, len(choices)))]
            tail = (O.OP_NORMALIZE, 0, 0, 0, 0, mean, std)
        elif r < 0.65:
            tail = (O.OP_DTC,)
            out_dt = int(rng.choice([O.DT_F32, O.DT_U8, O.DT_I16, O.DT_I32]))   # u8 -> i8 overflows above 127: undefined
        elif r < 0.68:
            out_dt = int(rng.choice([O.DT_F32, O.DT_I8]))   # no converting operator: the reference rejects the type
    if tail:
        ops.append(tail)
    n = int(rng.choice([1, 1, 1, 2, 3, 5]))
    resident = bool(rng.random() < 0.5)
    return dict(seed=seed, fmt=fmt, W=W, H=H, ops=ops, mode=mode, out_dt=out_dt, n=n, resident=resident, tail=tail)


def describe(c):
    ops = " -> ".join("%s%s" % (OP_NAMES[o[0]], tuple(o[1:]) if len(o) > 1 else "") for o in c["ops"]) or "(AutoConvert)"
    mem = c.get("mem") or ("resident" if c["resident"] else "host")
    if mem.startswith("padded"):
        mem += "%s" % (c["pads"],)
    if mem == "pinned" and c.get("misalign"):
        mem += "+%d" % c["misalign"]
    return "seed %d: %s %dx%d x%d %s %s | %s | out dt %d" % (c["seed"], FMT_NAMES[c["fmt"]], c["W"], c["H"], c["n"], mem,
                                                             c["mode"], ops, c["out_dt"])


def walk_shape(c, frame):
    """Final (w, h, format) of the image operators, by the oracle's stand-alone operators; None if one rejects."""
    cur, w, h, f = frame, c["W"], c["H"], c["fmt"]
    for op in c["ops"]:
        if op[0] in (O.OP_NORMALIZE, O.OP_DTC):
            break
        st, cur, info = O.run_op(op, cur, w, h, f)
        if st != 0:
            return None, None
        w, h, f = info
    return (w, h, f), cur


def arith_shape(c):
    """The output shape the chain's arguments ask for, whether or not an operator will reject them: the tensor handed to
    the processor has THIS shape (a mismatching one is undefined behaviour in the reference: ColorSpaceConvert validates
    formats only, image_operator.cc:367-399, and libyuv then writes input-sized rows into it)."""
    w, h, f = c["W"], c["H"], c["fmt"]
    for op in c["ops"]:
        if op[0] == O.OP_CROP:
            w, h = op[3] - op[1] + 1, op[4] - op[2] + 1
        elif op[0] == O.OP_ROTATE and op[1] in (90, 270):
            w, h = h, w
        elif op[0] == O.OP_RESIZE:
            w, h = op[1], op[2]
        elif op[0] == O.OP_CSC:
            f = op[1]
    if w <= 0 or h <= 0:
        return None
    return w, h, f


def oracle_side(seed):
    """Everything the checker computes for one case (runs in the worker process: a chain the reference would crash on
    takes the worker down, not the run)."""
    import gpu_util as G
    c = make_case(seed)
    frames = G.rand_frames(O, c["n"], c["W"], c["H"], c["fmt"], seed=seed * 7 + 1)
    shape, final = walk_shape(c, frames[0])
    r = dict(shape=shape, final=final)
    if c["mode"] == "tensor":
        a = arith_shape(c)
        if a is None:
            r["skip"] = "degenerate shape"
            return r
        w, h, f = a
        if w < 2 or h < 2:
            r["skip"] = "a tensor dimension of 1 is squeezed away (buffer.cc:106-152): raw output, undefined in the reference"
            return r
        r["dims"] = (1, h, w, 3)
        r["st"], r["want"] = G.oracle_chain(O, c["ops"] or [(O.OP_AUTO,)], frames, c["W"], c["H"], c["fmt"], c["out_dt"], r["dims"])
    return r


class OracleWorker:
    """A child interpreter answering oracle_side(seed) over pipes; restarted when it dies."""

    def __init__(self):
        self.p = None

    def start(self):
        import subprocess
        self.p = subprocess.Popen([sys.executable, os.path.abspath(__file__), "--oracle-worker"], stdin=subprocess.PIPE,
                                  stdout=subprocess.PIPE)

    def ask(self, seed):
        """One retry in a fresh worker: an earlier undefined-behaviour chain may have damaged the old worker's heap, and
        the crash then lands on an innocent seed."""
        r = self._ask_once(seed)
        if "crashed" in r:
            r = self._ask_once(seed)
        return r

    def _ask_once(self, seed):
        import pickle
        import struct
        if self.p is None or self.p.poll() is not None:
            self.start()
        try:
            self.p.stdin.write(struct.pack("<q", seed))
            self.p.stdin.flush()
            head = self.p.stdout.read(8)
            if len(head) < 8:
                raise EOFError
            n = struct.unpack("<q", head)[0]
            return pickle.loads(self.p.stdout.read(n))
        except (EOFError, BrokenPipeError, OSError):
            self.p.wait()
            rc = self.p.returncode
            self.p = None
            return dict(crashed=rc)

    def close(self):
        if self.p is not None and self.p.poll() is None:
            self.p.stdin.close()
            self.p.wait()


def oracle_worker_loop():
    import pickle
    import struct
    inp, out = sys.stdin.buffer, sys.stdout.buffer
    while True:
        head = inp.read(8)
        if len(head) < 8:
            return
        r = oracle_side(struct.unpack("<q", head)[0])
        blob = pickle.dumps(r, protocol=4)
        out.write(struct.pack("<q", len(blob)))
        out.write(blob)
        out.flush()


def _padded_layout(frame, W, H, fmt, y_pad, c_pad):
    """The same pixels with padded rows: -> (backing array, y_off, u_off, v_off, y_stride, c_stride, pixel_stride)."""
    cw, ch = (W + 1) // 2, (H + 1) // 2
    ys = W + y_pad
    y = frame[:W * H].reshape(H, W)
    if fmt in (O.FMT_NV12, O.FMT_NV21):
        cs = 2 * cw + c_pad
        back = np.full(H * ys + ch * cs + 64, 0xEE, dtype=np.uint8)
        back[:H * ys].reshape(H, ys)[:, :W] = y
        back[H * ys:H * ys + ch * cs].reshape(ch, cs)[:, :2 * cw] = frame[W * H:].reshape(ch, 2 * cw)
        first, second = H * ys, H * ys + 1
        u_off, v_off = (first, second) if fmt == O.FMT_NV12 else (second, first)
        return back, 0, u_off, v_off, ys, cs, 2
    cs = cw + c_pad
    back = np.full(H * ys + 2 * ch * cs + 64, 0xEE, dtype=np.uint8)
    back[:H * ys].reshape(H, ys)[:, :W] = y
    p1 = frame[W * H:W * H + cw * ch].reshape(ch, cw)
    p2 = frame[W * H + cw * ch:].reshape(ch, cw)
    back[H * ys:H * ys + ch * cs].reshape(ch, cs)[:, :cw] = p1
    back[H * ys + ch * cs:H * ys + 2 * ch * cs].reshape(ch, cs)[:, :cw] = p2
    o1, o2 = H * ys, H * ys + ch * cs
    u_off, v_off = (o2, o1) if fmt == O.FMT_YV12 else (o1, o2)   # YV12: Y, V, U; YV21: Y, U, V (buffer.cc:48-78)
    return back, 0, u_off, v_off, ys, cs, 1


def gpu_tensor_chain(c, frames, dims):
    """The product side of a tensor-mode case for every memory kind -> (status, ndarray or None)."""
    import ctypes as C
    import gpu_util as G
    from band_b200 import band_c as B
    mem = c.get("mem") or ("resident" if c["resident"] else "host")
    ops = c["ops"] or [(O.OP_AUTO,)]
    if mem in ("host", "resident"):
        return G.gpu_chain(ops, frames, c["W"], c["H"], c["fmt"], c["out_dt"], dims, device_resident=(mem == "resident"))
    L = B.lib()
    n, fb = frames.shape
    W, H, fmt = c["W"], c["H"], c["fmt"]
    device = mem.endswith("resident")
    proc = B.ImageProcessor(G.to_capi_ops(ops))
    tens = [B.Tensor(G.NP_OF_ORC[c["out_dt"]], dims, device=device) for _ in range(n)]
    bufs, host, dev, keep = [], None, None, []
    try:
        if mem == "pinned":
            pitch = (fb + 255) // 256 * 256
            host = L.BandHostAlloc(pitch * n + 64)
            for i in range(n):
                ptr = host + c.get("misalign", 0) + i * pitch
                C.memmove(ptr, frames[i].ctypes.data, fb)
                bufs.append(B.Buffer().set_raw(ptr, W, H, fmt))
        else:
            y_pad, c_pad = c["pads"]
            lays = [_padded_layout(frames[i], W, H, fmt, y_pad, c_pad) for i in range(n)]
            size = lays[0][0].size
            pitch = (size + 255) // 256 * 256
            if device:
                dev = L.BandDeviceAlloc(pitch * n)
                for i, lay in enumerate(lays):
                    if L.BandCopyToDevice(dev + i * pitch, lay[0].ctypes.data, size) != B.kBandOk:
                        raise B.BandError("device upload failed")
            for i, (back, yo, uo, vo, ys, cs, ps) in enumerate(lays):
                base = dev + i * pitch if device else back.ctypes.data
                keep.append(back)
                bufs.append(B.Buffer().set_yuv(base + yo, base + uo, base + vo, W, H, ys, cs, ps, fmt, device=device))
        st = proc.process_batch(bufs, tens) if n != 1 else proc.process(bufs[0], tens[0])
        out = np.stack([t.numpy() for t in tens]) if st == B.kBandOk else None
        return st, out
    finally:
        for x in bufs + tens + [proc]:
            x.close()
        if host:
            L.BandHostFree(host)
        if dev:
            L.BandDeviceFree(dev)


def run_case(c, worker, verbose=False):
    """-> (verdict, detail): 'same', 'both_reject', 'skipped', 'reference_crashes', or 'FAIL' with the difference."""
    import gpu_util as G
    from band_b200 import band_c as B
    frames = G.rand_frames(O, c["n"], c["W"], c["H"], c["fmt"], seed=c["seed"] * 7 + 1)
    r = worker.ask(c["seed"])
    if "crashed" in r:
        return "reference_crashes", "oracle worker died with %s" % r["crashed"]
    shape, final = r["shape"], r["final"]
    if c["mode"] == "image":
        if shape is None or c["tail"]:
            return "skipped", "image mode needs a chain the oracle walks"
        if not c["ops"]:
            return "skipped", "empty image chain"
        w, h, f = shape
        st, got, info = G.gpu_image_chain(c["ops"], frames[0], c["W"], c["H"], c["fmt"], f, O.DT_U8, w, h)
        if st != 0:
            return "FAIL", "oracle ok (%dx%d %s), product status %d: %s" % (w, h, FMT_NAMES[f], st, B.last_error())
        if info[:3] != (w, h, f):
            return "FAIL", "output info %s, oracle %s" % (info[:3], (w, h, f))
        bad = int((np.asarray(got) != np.asarray(final)).sum()) if got.shape == final.shape else -1
        return ("same", "") if bad == 0 else ("FAIL", "%d of %d bytes differ (image %dx%d %s)" % (bad, final.size, w, h, FMT_NAMES[f]))
    # tensor mode: the whole chain on both sides, the oracle's verdict is the reference's
    if "skip" in r:
        return "skipped", r["skip"]
    dims, st_o, want = r["dims"], r["st"], r["want"]
    st_g, got = gpu_tensor_chain(c, frames, dims)
    if verbose:
        print("  dims", dims, "oracle status", st_o, "product status", st_g, B.last_error() if st_g else "")
    if st_o == 0 and st_g != 0 and "input and output buffer formats must match" in B.last_error():
        # deliberate deviation: Crop / Flip / Rotate of an RGBA image into an RGB buffer passes the reference's
        # compatibility test (buffer.cc IsBufferFormatCompatible) and then moves 4-byte pixels into 3-byte rows -- past
        # the end of the output (larger frames abort the oracle worker with a corrupted heap). The product refuses.
        return "reference_overruns", ""
    if (st_o == 0) != (st_g == 0):
        return "FAIL", "oracle status %d, product status %d (%s), tensor dims %s" % (st_o, st_g, B.last_error(), dims)
    if st_o != 0:
        return "both_reject", ""
    g = np.ascontiguousarray(got).view(np.uint8)
    wv = np.ascontiguousarray(want).view(np.uint8)
    if g.shape != wv.shape:
        return "FAIL", "shape %s vs %s" % (got.shape, want.shape)
    bad = int((g != wv).sum())
    return ("same", "") if bad == 0 else ("FAIL", "%d of %d bytes differ (tensor %s)" % (bad, g.size, dims))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", type=int, default=500)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--only", type=int, default=None)
    ap.add_argument("--max-fail", type=int, default=40)
    ap.add_argument("--seeds", type=str, default=None, help="comma-separated seeds instead of a range")
    ap.add_argument("--quiet", action="store_true")
    ap.add_argument("--oracle-worker", action="store_true", help=argparse.SUPPRESS)
    a = ap.parse_args()
    if a.oracle_worker:
        oracle_worker_loop()
        return 0
    worker = OracleWorker()
    if a.only is not None:
        c = make_case(a.only)
        print(describe(c))
        print(run_case(c, worker, verbose=True))
        worker.close()
        return 0
    counts = {}
    fails = []
    t0 = time.time()
    seeds = [int(x) for x in a.seeds.split(",")] if a.seeds else range(a.seed, a.seed + a.cases)
    for s in seeds:
        c = make_case(s)
        try:
            v, detail = run_case(c, worker)
        except Exception as e:  # noqa: BLE001
            v, detail = "FAIL", "exception %s: %s" % (type(e).__name__, e)
        counts[v] = counts.get(v, 0) + 1
        if v == "reference_crashes":
            print("note", describe(c), "::", detail, flush=True)
        if v == "FAIL":
            fails.append(s)
            if not a.quiet:
                print("FAIL", describe(c), "::", detail, flush=True)
            if len(fails) >= a.max_fail:
                break
    worker.close()
    print("fuzz_chains: seeds %d..%d in %.0f s: %s; failing seeds: %s" % (a.seed, a.seed + a.cases - 1, time.time() - t0, counts, fails))
    return 1 if fails else 0


if __name__ == "__main__":
    sys.exit(main())
