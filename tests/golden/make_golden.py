"""Generates the committed fixtures under tests/golden/ (run in the build container only:
it reads /root/reference, which does not exist on the GPU box).

  hippo.npz           the reference's own image fixtures (band/test/data/hippo.jpg,
                      hippo_crop_256.jpg, hippo_resize_256.jpg; used by
                      band/test/image_operation_test.cc:31-44,143-159) decoded to RGB arrays.
                      Decoder: PIL/libjpeg (the reference decodes with stb_image,
                      band/test/image_util.cc:26-38; both are exact-IDCT-class decoders and the
                      crop fixture decodes byte-identical to the crop of hippo.jpg with PIL).
  oracle_frozen.json  sha256 digests of the CPU oracle's outputs on seeded inputs for every
                      BASELINE.json chain shape (small sizes). Where no reference test pins the
                      values (YUV->RGB, Normalize, rotate direction, scaler special cases) the
                      frozen oracle IS the spec; this file detects any later drift of it.

Usage:  python tests/golden/make_golden.py
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF_DATA = "/root/reference/band/test/data"


def make_hippo():
    from PIL import Image
    arrs = {}
    for key, name in (("hippo", "hippo.jpg"), ("crop_256", "hippo_crop_256.jpg"), ("resize_256", "hippo_resize_256.jpg")):
        arrs[key] = np.asarray(Image.open(os.path.join(REF_DATA, name)).convert("RGB"), dtype=np.uint8)
    np.savez_compressed(os.path.join(HERE, "hippo.npz"), **arrs)
    for k, v in arrs.items():
        print(k, v.shape)


def frozen_cases():
    """(name, ops, W, H, fmt, out_dtype, out_dims, seed). Shared with tests/test_oracle_golden.py."""
    from oracle import oracle as O
    N = (O.OP_NORMALIZE, 0, 0, 0, 0, 127.5, 127.5)
    cases = []
    for fmt_name, fmt in (("nv12", O.FMT_NV12), ("nv21", O.FMT_NV21), ("yv12", O.FMT_YV12), ("yv21", O.FMT_YV21)):
        cases.append(("c1_%s_64x48_to_22x22_f32" % fmt_name, [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, 22, 22), N],
                      64, 48, fmt, O.DT_F32, (1, 22, 22, 3), 11))
        cases.append(("c2_%s_96x54_to_30x30_u8" % fmt_name, [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, 30, 30)],
                      96, 54, fmt, O.DT_U8, (1, 30, 30, 3), 12))
        cases.append(("odd_%s_65x47_to_23x21_u8" % fmt_name, [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, 23, 21)],
                      65, 47, fmt, O.DT_U8, (1, 21, 23, 3), 13))
    # C3 shape: Rotate 90 -> centre Crop -> CSC -> Resize -> Normalize on planar 4:2:0 (192x108 stands for 3840x2160)
    cases.append(("c3_yv12_192x108_rot90_crop_to_26x26_f32",
                  [(O.OP_ROTATE, 90), (O.OP_CROP, 0, 42, 107, 149), (O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, 26, 26), N],
                  192, 108, O.FMT_YV12, O.DT_F32, (1, 26, 26, 3), 14))
    # C5 shapes: RGBA -> RGB / gray with power-of-two and general ratios, up and down
    for s, d in ((64, 16), (64, 32), (64, 48), (64, 24), (32, 64), (32, 80), (48, 16)):
        cases.append(("c5_rgba_%d_to_rgb_%d" % (s, d), [(O.OP_AUTO,)], s, s, O.FMT_RGBA, O.DT_U8, (1, d, d, 3), 15))
        cases.append(("c5_rgba_%d_to_gray_%d" % (s, d), [(O.OP_CSC, O.FMT_GRAY), (O.OP_RESIZE, d, d)], s, s,
                      O.FMT_RGBA, O.DT_U8, (d, d), 16))
    return cases


def run_case(case):
    from oracle import oracle as O
    name, ops, W, H, fmt, odt, odims, seed = case
    rng = np.random.default_rng(seed)
    frame = rng.integers(0, 256, size=O.frame_bytes(W, H, fmt), dtype=np.uint8)
    if len(odims) == 2:  # gray image output: use a [1,h,w,1]-less raw tensor is not an image; run ops standalone
        cur, w, h, f = frame, W, H, fmt
        for op in ops:
            st, cur, (w, h, f) = O.run_op(op, cur, w, h, f)
            assert st == 0, (name, st)
        return cur
    st, out = O.run_chain(ops, frame, W, H, fmt, odt, odims)
    assert st == 0, (name, st)
    return out


def make_frozen():
    digests = {}
    for case in frozen_cases():
        out = run_case(case)
        digests[case[0]] = hashlib.sha256(np.ascontiguousarray(out).tobytes()).hexdigest()
    with open(os.path.join(HERE, "oracle_frozen.json"), "w") as f:
        json.dump(digests, f, indent=1, sort_keys=True)
    print("froze %d oracle cases" % len(digests))


if __name__ == "__main__":
    make_hippo()
    make_frozen()
