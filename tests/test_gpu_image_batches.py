"""Batches whose outputs are IMAGES (gray, 4:2:0), through BandImageProcessorProcessBuffers: every frame against the oracle's
stand-alone operators. Two layouts of the same batch: frames carved from one allocation at a uniform pitch (the shim's strided
mode) and frames at irregular offsets / in separate allocations (its pointer-table mode) -- the plane kernels, the 4:2:0
index-map kernels and the RGB -> 4:2:0 kernels all take both."""
import ctypes as C

import numpy as np
import pytest

from gpu_util import assert_same_bits, rand_frames

pytestmark = pytest.mark.gpu


def _oracle(O, ops, fr, W, H, fmt):
    cur, w, h, f = fr, W, H, fmt
    for op in ops:
        st, cur, (w, h, f) = O.run_op(op, cur, w, h, f)
        assert st == 0, (op, st)
    return cur, w, h, f


CHAINS = [
    ("NV21", 640, 360, [("RESIZE", 426, 240)]),                     # plane-rows kernels (Y + interleaved chroma)
    ("YV12", 640, 360, [("RESIZE", 320, 180)]),                     # 2x2 box, three planes
    ("NV12", 320, 180, [("RESIZE", 640, 360)]),                     # exact 2x up-samplers
    ("YV21", 322, 182, [("RESIZE", 215, 97)]),                      # odd strides: the fallbacks
    ("NV21", 640, 360, [("ROTATE", 90)]),
    ("YV12", 640, 360, [("CROP", 32, 16, 351, 255), ("FLIP", 1, 0)]),
    ("RGBA", 320, 180, [("CSC", "NV21")]),
    ("RGB", 320, 180, [("CSC", "GRAY"), ("RESIZE", 640, 360)]),     # gray conversion + exact 2x
    ("NV12", 640, 360, [("CSC", "GRAY")]),
    ("GRAY", 640, 360, [("RESIZE", 400, 200)]),
]


@pytest.mark.parametrize("layout", ["uniform", "table"])
@pytest.mark.parametrize("fmt_name,W,H,chain", CHAINS)
def test_image_output_batches(O, fmt_name, W, H, chain, layout):
    from band_b200 import band_c as B
    L = B.lib()
    fmt = getattr(O, "FMT_" + fmt_name)
    kinds = {"RESIZE": O.OP_RESIZE, "ROTATE": O.OP_ROTATE, "CROP": O.OP_CROP, "FLIP": O.OP_FLIP, "CSC": O.OP_CSC}
    ops = []
    for op in chain:
        args = [getattr(O, "FMT_" + a) if isinstance(a, str) else a for a in op[1:]]
        ops.append((kinds[op[0]],) + tuple(args))
    n = 5
    frames = rand_frames(O, n, W, H, fmt, seed=W + len(chain) * 13)
    wants = []
    for f in frames:
        want, ow, oh, ofmt = _oracle(O, ops, f, W, H, fmt)
        wants.append(want)
    fb, ob = frames.shape[1], wants[0].size
    pitch_in, pitch_out = (fb + 255) // 256 * 256, (ob + 255) // 256 * 256
    # "table": frame i sits at an irregular, still word-aligned offset -> no common stride between consecutive frames
    skew = [0, 0, 0, 0, 0] if layout == "uniform" else [0, 256, 64, 1024, 16]
    din, dout = L.BandDeviceAlloc((pitch_in + 2048) * n), L.BandDeviceAlloc((pitch_out + 2048) * n)
    assert din and dout
    ins, outs = [], []
    try:
        for i in range(n):
            pi = din + i * (pitch_in + (0 if layout == "uniform" else 2048)) + skew[i]
            po = dout + i * (pitch_out + (0 if layout == "uniform" else 2048)) + skew[(i + 2) % n]
            assert L.BandCopyToDevice(pi, frames[i].ctypes.data, fb) == B.kBandOk
            ins.append(B.Buffer().set_raw(pi, W, H, fmt, device=True))
            outs.append((po, B.Buffer().set_raw(po, ow, oh, ofmt, device=True)))
        from gpu_util import to_capi_ops
        proc = B.ImageProcessor(to_capi_ops(ops))
        ia = (C.c_void_p * n)(*[b.handle for b in ins])
        oa = (C.c_void_p * n)(*[b.handle for _, b in outs])
        for rnd in range(2):  # the second call takes the repeat-call path
            assert L.BandImageProcessorProcessBuffers(proc.handle, ia, oa, n, 0) == B.kBandOk, B.last_error()
            for i in range(n):
                got = np.empty(ob, dtype=np.uint8)
                assert L.BandCopyToHost(got.ctypes.data, outs[i][0], ob) == B.kBandOk
                assert_same_bits(got, wants[i], "%s %s frame %d (%s, call %d)" % (fmt_name, chain, i, layout, rnd))
        proc.close()
    finally:
        for b in ins:
            b.close()
        for _, b in outs:
            b.close()
        L.BandDeviceFree(din)
        L.BandDeviceFree(dout)
