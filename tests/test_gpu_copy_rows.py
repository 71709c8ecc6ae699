"""Same-size conversions (ColorSpaceConvert without a Resize, optionally next to a Crop): the word-access kernels of
band_b200/csrc/cuda/bandcu_copy.cu against the oracle, bit for bit, next to the shapes that must fall back to the
pixel-per-thread kernel (widths that are not a multiple of four, odd crop origins after the conversion).

Reference: LibyuvImageOperator::ColorSpaceConvert (libyuv_image_operator.cc:1709-1727) and ::Crop (:1609-1625)."""
import numpy as np
import pytest

from gpu_util import assert_same_bits, gpu_image_chain, gpu_op, rand_frames

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def O():
    from oracle import oracle
    return oracle


def _oracle_ops(O, ops, fr, W, H, fmt):
    cur, w, h, f = fr, W, H, fmt
    for op in ops:
        st, cur, (w, h, f) = O.run_op(op, cur, w, h, f)
        assert st == 0, (op, st)
    return cur, w, h, f


YUV = ["NV21", "NV12", "YV12", "YV21"]
SIZES = [(64, 48), (68, 35), (36, 22), (34, 22), (1920, 1080), (4, 2), (8, 1)]


@pytest.mark.parametrize("dst_name", ["RGB", "RGBA", "GRAY"])
@pytest.mark.parametrize("fmt_name", YUV)
@pytest.mark.parametrize("W,H", SIZES)
def test_yuv_to_packed_same_size(O, fmt_name, dst_name, W, H):
    fmt, dst = getattr(O, "FMT_" + fmt_name), getattr(O, "FMT_" + dst_name)
    fr = rand_frames(O, 1, W, H, fmt, 900 + W + H)[0]
    st, want, winfo = O.run_op((O.OP_CSC, dst), fr, W, H, fmt)
    st2, got, info = gpu_op((O.OP_CSC, dst), fr, W, H, fmt)
    assert st == st2
    if st == 0:
        assert info[:3] == tuple(winfo)
        assert_same_bits(got, want, "%s -> %s %dx%d" % (fmt_name, dst_name, W, H))


@pytest.mark.parametrize("src_name,dst_name", [("RGBA", "RGB"), ("RGBA", "GRAY"), ("RGB", "GRAY")])
@pytest.mark.parametrize("W,H", [(64, 48), (20, 31), (33, 21), (1920, 1080), (4, 1)])
def test_packed_to_packed_same_size(O, src_name, dst_name, W, H):
    fmt, dst = getattr(O, "FMT_" + src_name), getattr(O, "FMT_" + dst_name)
    fr = rand_frames(O, 1, W, H, fmt, 950 + W + H)[0]
    st, want, winfo = O.run_op((O.OP_CSC, dst), fr, W, H, fmt)
    st2, got, info = gpu_op((O.OP_CSC, dst), fr, W, H, fmt)
    assert st == 0 and st2 == 0 and info[:3] == tuple(winfo)
    assert_same_bits(got, want, "%s -> %s %dx%d" % (src_name, dst_name, W, H))


@pytest.mark.parametrize("fmt_name", YUV)
@pytest.mark.parametrize("box", [(3, 5, 42, 24), (4, 2, 43, 21), (8, 4, 71, 51), (1, 1, 38, 30), (0, 0, 99, 59)])
@pytest.mark.parametrize("dst_name", ["RGB", "RGBA"])
def test_crop_then_convert(O, fmt_name, box, dst_name):
    """Crop in the YUV domain (CropNv / CropYv take the chroma origin at origin/2, :950-1037), then convert."""
    W, H = 100, 60
    fmt, dst = getattr(O, "FMT_" + fmt_name), getattr(O, "FMT_" + dst_name)
    fr = rand_frames(O, 1, W, H, fmt, 970)[0]
    ops = [(O.OP_CROP,) + box, (O.OP_CSC, dst)]
    want, w, h, f = _oracle_ops(O, ops, fr, W, H, fmt)
    st, got, info = gpu_image_chain(ops, fr, W, H, fmt, dst, O.DT_U8, w, h)
    assert st == 0 and info[:3] == (w, h, dst)
    assert_same_bits(got, want, "crop %s -> %s %s" % (fmt_name, dst_name, box))


@pytest.mark.parametrize("fmt_name", ["NV21", "YV12"])
@pytest.mark.parametrize("box", [(4, 2, 43, 21), (3, 5, 42, 24), (8, 4, 71, 51), (2, 3, 41, 30)])
def test_convert_then_crop(O, fmt_name, box):
    """Convert, then crop the RGB image: even origins fold into the word-access kernel, odd ones do not."""
    W, H = 100, 60
    fmt = getattr(O, "FMT_" + fmt_name)
    fr = rand_frames(O, 1, W, H, fmt, 980)[0]
    ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_CROP,) + box]
    want, w, h, f = _oracle_ops(O, ops, fr, W, H, fmt)
    st, got, info = gpu_image_chain(ops, fr, W, H, fmt, O.FMT_RGB, O.DT_U8, w, h)
    assert st == 0 and info[:3] == (w, h, O.FMT_RGB)
    assert_same_bits(got, want, "%s -> RGB, crop %s" % (fmt_name, box))


def test_every_luma_value_against_every_chroma_pair_sample(O):
    """256 luma values x a 64 x 64 grid of (U, V): one 1024 x 1024 NV12 frame through the word-access kernel."""
    W = H = 1024
    y = np.tile(np.arange(256, dtype=np.uint8), (H, W // 256))
    uu, vv = np.meshgrid(np.arange(0, 256, 4, dtype=np.uint8), np.arange(0, 256, 4, dtype=np.uint8))
    uv = np.zeros((H // 2, W // 2, 2), dtype=np.uint8)
    uv[..., 0] = np.tile(uu, (8, 8))
    uv[..., 1] = np.tile(vv, (8, 8))
    fr = np.concatenate([y.reshape(-1), uv.reshape(-1)])
    st, want, _ = O.run_op((O.OP_CSC, O.FMT_RGB), fr, W, H, O.FMT_NV12)
    st2, got, _ = gpu_op((O.OP_CSC, O.FMT_RGB), fr, W, H, O.FMT_NV12)
    assert st == 0 and st2 == 0
    assert_same_bits(got, want, "NV12 value grid")
