"""Row-form kernels of round 2 against the oracle, bit for bit, at sizes that take the vector path (whole words / vectors,
aligned planes), sizes that fall back to the element-wise kernels, and sizes whose last tile is partial:

  * RGB / RGBA -> 4:2:0 (rgb_to_yuv420_rows_kernel; RAWToI420 / ABGRToNV12 / NV21 / I420, libyuv_image_operator.cc:523-674)
  * 4:2:0 <-> 4:2:0 chroma repacks (yuv_repack_rows_kernel; :125-162, :213-250, :302-367)
  * quarter-turns as word transposes (bandcu_turn.cu; Rotate, :697-887) on every format
"""
import pytest

from gpu_util import assert_same_bits, gpu_chain, gpu_op, oracle_chain, rand_frames

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def O():
    from oracle import oracle
    return oracle


@pytest.mark.parametrize("W,H", [(640, 480), (644, 482), (1924, 1082), (322, 242), (321, 241), (8, 2), (4, 2)])
@pytest.mark.parametrize("src", ["RGB", "RGBA"])
@pytest.mark.parametrize("dst", ["NV12", "NV21", "YV12", "YV21"])
def test_rgb_to_yuv420(O, W, H, src, dst):
    fmt = getattr(O, "FMT_" + src)
    fr = rand_frames(O, 1, W, H, fmt, 2100 + W)[0]
    op = (O.OP_CSC, getattr(O, "FMT_" + dst))
    st, want, winfo = O.run_op(op, fr, W, H, fmt)
    st2, got, info = gpu_op(op, fr, W, H, fmt)
    assert st == 0 and st2 == 0
    assert info[:3] == tuple(winfo)
    assert_same_bits(got, want, "%s -> %s %dx%d" % (src, dst, W, H))


@pytest.mark.parametrize("W,H", [(640, 480), (1920, 1080), (656, 482), (648, 480), (322, 242), (321, 241), (16, 2)])
@pytest.mark.parametrize("src,dst", [("NV12", "YV12"), ("NV12", "YV21"), ("NV21", "YV12"), ("YV12", "NV12"), ("YV21", "NV21"),
                                     ("YV12", "NV21"), ("NV12", "NV21"), ("NV21", "NV12"), ("YV12", "YV21")])
def test_yuv_repacks(O, W, H, src, dst):
    fmt = getattr(O, "FMT_" + src)
    fr = rand_frames(O, 1, W, H, fmt, 2200 + W)[0]
    op = (O.OP_CSC, getattr(O, "FMT_" + dst))
    st, want, winfo = O.run_op(op, fr, W, H, fmt)
    st2, got, info = gpu_op(op, fr, W, H, fmt)
    assert st == 0 and st2 == 0
    assert info[:3] == tuple(winfo)
    assert_same_bits(got, want, "%s -> %s %dx%d" % (src, dst, W, H))


@pytest.mark.parametrize("angle", [90, 270])
@pytest.mark.parametrize("fmt_name", ["GRAY", "RGB", "RGBA", "NV21", "NV12", "YV12", "YV21"])
@pytest.mark.parametrize("W,H", [(644, 484), (256, 128), (132, 36), (1284, 724), (8, 8), (4, 4)])
def test_quarter_turn_word_path_and_partial_tiles(O, angle, fmt_name, W, H):
    fmt = getattr(O, "FMT_" + fmt_name)
    fr = rand_frames(O, 1, W, H, fmt, 2300 + W)[0]
    st, want, winfo = O.run_op((O.OP_ROTATE, angle), fr, W, H, fmt)
    st2, got, info = gpu_op((O.OP_ROTATE, angle), fr, W, H, fmt)
    assert st == 0 and st2 == 0
    assert info[:3] == tuple(winfo)
    assert_same_bits(got, want, "rotate %d %s %dx%d" % (angle, fmt_name, W, H))


@pytest.mark.parametrize("W,H", [(640, 480), (644, 482), (1284, 724), (322, 242), (8, 2)])
@pytest.mark.parametrize("fmt_name", ["NV12", "NV21", "YV12", "YV21"])
@pytest.mark.parametrize("resident", [False, True])
def test_convert_then_normalize_without_resize(O, W, H, fmt_name, resident):
    """ColorSpaceConvert(RGB) -> Normalize(127.5, 127.5) into a float32 tensor at frame size (yuv_copy_rows_f32_kernel when
    the width is a multiple of four and the tensor is 16-byte aligned; the pixel-per-thread kernel otherwise)."""
    fmt = getattr(O, "FMT_" + fmt_name)
    fr = rand_frames(O, 2, W, H, fmt, 2400 + W)
    ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_NORMALIZE, 0, 0, 0, 0, 127.5, 127.5)]
    st, want = oracle_chain(O, ops, fr, W, H, fmt, O.DT_F32, (1, H, W, 3))
    st2, got = gpu_chain(ops, fr, W, H, fmt, O.DT_F32, (1, H, W, 3), device_resident=resident)
    assert st == 0 and st2 == 0
    assert_same_bits(got, want, "%s %dx%d -> RGB f32 (no resize)" % (fmt_name, W, H))
