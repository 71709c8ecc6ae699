"""CPU tests: the oracle against every golden vector / known-answer the reference's own tests hold for
this path (/root/reference/band/test/image_operation_test.cc), plus the frozen
digests of tests/golden/oracle_frozen.json. No GPU needed."""
import hashlib
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN


def test_crop_simple_kat(O):
    """image_operation_test.cc:12-29: 3x3 gray {1..9}, Crop(0,0,1,1) -> {1,2,4,5}."""
    st, out, (w, h, f) = O.run_op((O.OP_CROP, 0, 0, 1, 1), np.arange(1, 10, dtype=np.uint8), 3, 3, O.FMT_GRAY)
    assert st == O.OK and (w, h, f) == (2, 2, O.FMT_GRAY)
    assert out.tolist() == [1, 2, 4, 5]


def test_crop_hippo_exact(O, hippo):
    """image_operation_test.cc:31-44: Crop(0,0,255,255) of hippo.jpg == hippo_crop_256.jpg (exact)."""
    src = hippo["hippo"]
    H, W = src.shape[:2]
    st, out, (w, h, f) = O.run_op((O.OP_CROP, 0, 0, 255, 255), src.reshape(-1), W, H, O.FMT_RGB)
    assert st == O.OK and (w, h, f) == (256, 256, O.FMT_RGB)
    assert np.array_equal(out.reshape(256, 256, 3), hippo["crop_256"])


@pytest.mark.parametrize("box", [(0, 0, 600, 400), (255, 0, 0, 255), (0, 255, 255, 0), (-1, -1, 256, 256),
                                 (0, 0, -1, -1), (-1, -1, -1, -1)])
def test_crop_failures(O, hippo, box):
    """image_operation_test.cc:46-75: every bad box is kInvalidArgument."""
    src = hippo["hippo"]
    H, W = src.shape[:2]
    st, _, _ = O.run_op((O.OP_CROP,) + box, src.reshape(-1), W, H, O.FMT_RGB)
    assert st == O.INVALID_ARGUMENT


def test_rgb_to_gray_within_one(O, hippo):
    """image_operation_test.cc:77-119: |gray - (0.299R+0.587G+0.114B)| <= 1."""
    src = hippo["hippo"]
    H, W = src.shape[:2]
    st, out, (w, h, f) = O.run_op((O.OP_CSC, O.FMT_GRAY), src.reshape(-1), W, H, O.FMT_RGB)
    assert st == O.OK and (w, h, f) == (W, H, O.FMT_GRAY)
    want = 0.299 * src[..., 0] + 0.587 * src[..., 1] + 0.114 * src[..., 2]
    assert np.abs(out.reshape(H, W).astype(np.float64) - want).max() <= 1.0


def test_rgb_to_gray_exhaustive_within_one(O):
    """The 8-bit J400 form satisfies the reference's +-1 assertion for EVERY RGB triple."""
    r, g, b = np.meshgrid(np.arange(256), np.arange(0, 256, 3), np.arange(256), indexing="ij")
    rgb = np.stack([r, g, b], -1).astype(np.uint8).reshape(-1, 3)
    n = rgb.shape[0]
    W = 4096
    H = n // W
    rgb = rgb[:W * H]
    st, out, _ = O.run_op((O.OP_CSC, O.FMT_GRAY), rgb.reshape(-1), W, H, O.FMT_RGB)
    assert st == O.OK
    want = 0.299 * rgb[:, 0] + 0.587 * rgb[:, 1] + 0.114 * rgb[:, 2]
    assert np.abs(out.astype(np.float64) - want).max() <= 1.0


def test_rotate_round_trip(O, hippo):
    """image_operation_test.cc:121-141: Rotate(90) swaps dims; Rotate(270) after it restores the image."""
    src = hippo["hippo"]
    H, W = src.shape[:2]
    st, r90, (w, h, f) = O.run_op((O.OP_ROTATE, 90), src.reshape(-1), W, H, O.FMT_RGB)
    assert st == O.OK and (w, h) == (H, W)
    # counter-clockwise (libyuv_image_operator.cc:697-708): dst[r][c] = src[c][W-1-r]
    assert np.array_equal(r90.reshape(W, H, 3), np.rot90(src, 1))
    st, back, (w2, h2, _) = O.run_op((O.OP_ROTATE, 270), r90, w, h, O.FMT_RGB)
    assert st == O.OK and (w2, h2) == (W, H)
    assert np.array_equal(back.reshape(H, W, 3), src)


def test_resize_hippo_fixture(O, hippo):
    """image_operation_test.cc:143-159: Resize(256,256) within 3 of hippo_resize_256.jpg on the first
    W*H bytes the reference compares; SURVEY 8c: the fixture also discriminates blender variants
    through the MEAN error (< 0.25 only for the truncating 7-bit ARGB column filter)."""
    src = hippo["hippo"]
    H, W = src.shape[:2]
    st, out, (w, h, f) = O.run_op((O.OP_RESIZE, 256, 256), src.reshape(-1), W, H, O.FMT_RGB)
    assert st == O.OK and (w, h, f) == (256, 256, O.FMT_RGB)
    got = out.astype(np.int32)
    want = hippo["resize_256"].reshape(-1).astype(np.int32)
    n = 256 * 256  # GetNumElements(): what the reference loop covers
    assert np.abs(got[:n] - want[:n]).max() <= 3
    diff = np.abs(got - want)
    assert diff.mean() < 0.25
    assert diff.max() <= 5  # JPEG re-encode noise of the fixture over the whole image


def test_normalize_range(O, hippo):
    """tfl_minimal_test.cc:504-520: Resize(224,224) + Normalize(127.5,127.5) lands in [-1,1]."""
    src = hippo["hippo"]
    H, W = src.shape[:2]
    ops = [(O.OP_RESIZE, 224, 224), (O.OP_NORMALIZE, 0, 0, 0, 0, 127.5, 127.5)]
    st, out = O.run_chain(ops, src.reshape(-1), W, H, O.FMT_RGB, O.DT_F32, (1, 224, 224, 3))
    assert st == O.OK
    assert out.min() >= -1.0 and out.max() <= 1.0
    st, u8 = O.run_chain(ops[:1], src.reshape(-1), W, H, O.FMT_RGB, O.DT_U8, (1, 224, 224, 3))
    assert st == O.OK
    want = (u8.astype(np.float32) - np.float32(127.5)) / np.float32(127.5)
    assert np.array_equal(out.view(np.uint32), want.view(np.uint32))


def test_yuv_pixel_anchors_and_float_model(O):
    """No reference test pins YUV->RGB (parity unpinned). Anchors from the BT.601 limited-range
    definition plus a float model of the same matrix: the fixed-point result stays within 1."""
    import ctypes as C
    L = O.lib()

    def px(y, u, v):
        r, g, b = C.c_uint8(), C.c_uint8(), C.c_uint8()
        L.orc_YuvPixel(y, u, v, C.byref(r), C.byref(g), C.byref(b))
        return r.value, g.value, b.value

    assert px(16, 128, 128) == (0, 0, 0)
    assert px(235, 128, 128) == (255, 255, 255)
    assert px(0, 128, 128) == (0, 0, 0) and px(255, 128, 128) == (255, 255, 255)
    rng = np.random.default_rng(1)
    for y, u, v in rng.integers(0, 256, size=(4000, 3)):
        yf = 1.164 * (int(y) - 16)
        # libyuv's 6-bit matrix: VR=102/64, UG=25/64, VG=52/64, UB capped at 128/64 (2.0, not 2.018)
        want = (yf + 1.59375 * (int(v) - 128), yf - 0.390625 * (int(u) - 128) - 0.8125 * (int(v) - 128),
                yf + 2.0 * (int(u) - 128))
        got = px(int(y), int(u), int(v))
        for a, b in zip(got, want):
            assert abs(a - min(255.0, max(0.0, b))) <= 1.01, (y, u, v, got, want)


def test_frozen_digests(O):
    """The oracle has not drifted from the digests frozen in tests/golden/oracle_frozen.json."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(GOLDEN, "make_golden.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    with open(os.path.join(GOLDEN, "oracle_frozen.json")) as f:
        frozen = json.load(f)
    cases = mg.frozen_cases()
    assert sorted(c[0] for c in cases) == sorted(frozen)
    for case in cases:
        out = mg.run_case(case)
        assert hashlib.sha256(np.ascontiguousarray(out).tobytes()).hexdigest() == frozen[case[0]], case[0]


def test_yuv_to_rgb_agrees_with_opencv_as_an_independent_witness(O):
    """K1 (YUV -> RGB values) is pinned by no vector the reference holds: the restated libyuv arithmetic IS
    the spec. As a second, independent witness OpenCV's BT.601 limited-range conversion (its own 20-bit fixed point, not
    libyuv's 6-bit one) must agree with the oracle within a few counts on random frames -- enough to catch a U/V swap, a
    channel-order mistake, a range (full vs limited) mistake or a wrong plane order in any of the four 4:2:0 layouts."""
    cv2 = pytest.importorskip("cv2")
    W, H = 64, 48
    rng = np.random.default_rng(99)
    cases = [(O.FMT_NV12, cv2.COLOR_YUV2RGB_NV12), (O.FMT_NV21, cv2.COLOR_YUV2RGB_NV21),
             (O.FMT_YV12, cv2.COLOR_YUV2RGB_YV12), (O.FMT_YV21, cv2.COLOR_YUV2RGB_I420)]
    for fmt, code in cases:
        # mid-range chroma keeps both converters away from their clamps, where rounding differences would be amplified
        n = O.frame_bytes(W, H, fmt)
        frame = rng.integers(16, 236, size=n, dtype=np.uint8)
        frame[W * H:] = rng.integers(64, 192, size=n - W * H, dtype=np.uint8)
        st, ours, (w, h, f) = O.run_op((O.OP_CSC, O.FMT_RGB), frame, W, H, fmt)
        assert st == 0 and (w, h, f) == (W, H, O.FMT_RGB)
        theirs = cv2.cvtColor(frame.reshape(H * 3 // 2, W), code)
        diff = np.abs(ours.reshape(H, W, 3).astype(np.int32) - theirs.astype(np.int32))
        assert diff.max() <= 3, (fmt, int(diff.max()))
        assert diff.mean() < 1.0, (fmt, float(diff.mean()))
        # the witness is sensitive: swapping the chroma planes moves the result far outside the tolerance
        swapped_code = {cv2.COLOR_YUV2RGB_NV12: cv2.COLOR_YUV2RGB_NV21, cv2.COLOR_YUV2RGB_NV21: cv2.COLOR_YUV2RGB_NV12,
                        cv2.COLOR_YUV2RGB_YV12: cv2.COLOR_YUV2RGB_I420, cv2.COLOR_YUV2RGB_I420: cv2.COLOR_YUV2RGB_YV12}[code]
        wrong = cv2.cvtColor(frame.reshape(H * 3 // 2, W), swapped_code)
        assert np.abs(ours.reshape(H, W, 3).astype(np.int32) - wrong.astype(np.int32)).max() > 20


def test_vector_rows_equal_the_scalar_restatement(O, tmp_path):
    """The oracle's AVX2 rows (4:2:0 -> RGB / RGBA, RGB24 -> ARGB, InterpolateRow: what makes the timed CPU baseline more than a
    scalar straw man) against its scalar restatement: (1) the conversion on all 2^24 (y, u, v) triples inside the library,
    (2) whole chains and the ...ToABGR rows against a second build of the same sources WITHOUT AVX2 (-march=x86-64), at
    sizes with and without row tails."""
    import ctypes as C
    import glob
    import os
    import shutil
    import subprocess
    L = O.lib()
    L.orc_yuv_rows_selftest.restype = C.c_long
    assert L.orc_yuv_rows_selftest() == 0
    here = os.path.dirname(os.path.abspath(O.__file__))
    cc = os.environ.get("CC") or shutil.which("gcc") or "gcc"
    so = str(tmp_path / "liborc_scalar.so")
    srcs = sorted(glob.glob(os.path.join(here, "*.c")))
    subprocess.check_call([cc, "-O2", "-march=x86-64", "-ffp-contract=off", "-fPIC", "-std=c11", "-shared", "-o", so] + srcs)
    S = C.CDLL(so)
    S.orc_run_batch.argtypes = L.orc_run_batch.argtypes
    S.orc_run_batch.restype = L.orc_run_batch.restype
    S.orc_yuv_rows_selftest.restype = C.c_long
    assert S.orc_yuv_rows_selftest() == 0      # (nothing to compare there: documents that this build has no vector rows)
    rng = np.random.default_rng(5)
    norm = (O.OP_NORMALIZE, 0, 0, 0, 0, 127.5, 127.5)
    for fmt in (O.FMT_NV12, O.FMT_NV21, O.FMT_YV12, O.FMT_YV21):
        for (W, H, dw, dh) in [(1920, 1080, 224, 224), (642, 362, 300, 200), (646, 360, 640, 352), (36, 22, 20, 11), (14, 10, 9, 7),
                               (320, 240, 500, 333)]:
            ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, dw, dh), norm]
            frames = rng.integers(0, 256, size=(2, O.frame_bytes(W, H, fmt)), dtype=np.uint8)
            descs = O.make_descs(ops)
            dims = (C.c_int * 4)(1, dh, dw, 3)
            outs = []
            for lib in (L, S):
                out = np.zeros((2, 1, dh, dw, 3), dtype=np.float32)
                st = lib.orc_run_batch(descs, len(ops), frames.ctypes.data, frames.shape[1], W, H, fmt, out.ctypes.data,
                                       out[0].nbytes, O.DT_F32, dims, 4, 2)
                assert st == 0
                outs.append(out)
            assert np.array_equal(outs[0].view(np.uint32), outs[1].view(np.uint32)), (fmt, W, H, dw, dh)
    # RGBA rows (…ToABGR): semi-planar in both orders
    vp, ci = C.c_void_p, C.c_int
    for name in ("orc_NV12ToABGR", "orc_NV21ToABGR"):
        for (W, H) in [(1920, 8), (50, 6), (16, 2), (15, 3)]:
            y = rng.integers(0, 256, size=W * H, dtype=np.uint8)
            uv = rng.integers(0, 256, size=2 * ((W + 1) // 2) * ((H + 1) // 2), dtype=np.uint8)
            res = []
            for lib in (L, S):
                f = getattr(lib, name)
                f.argtypes = [vp, ci, vp, ci, vp, ci, ci, ci]
                out = np.zeros(W * H * 4, dtype=np.uint8)
                assert f(y.ctypes.data, W, uv.ctypes.data, 2 * ((W + 1) // 2), out.ctypes.data, W * 4, W, H) == 0
                res.append(out)
            assert np.array_equal(res[0], res[1]), (name, W, H)
