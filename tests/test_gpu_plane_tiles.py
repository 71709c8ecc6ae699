"""Crop / Rotate / Flip with nothing else fused, at sizes that span several 32-row tiles of the tile-staged plane copy
(band_b200/csrc/cuda/bandcu_generic.cu: plane_map_tile_kernel; 1 - 4 bytes per element, 4:2:0 planes and packed images)
against the oracle, bit for bit. Reference: LibyuvImageOperator::Rotate / ::Crop / ::Flip*
(libyuv_image_operator.cc:1609-1707)."""
import pytest

from gpu_util import assert_same_bits, gpu_image_chain, gpu_op, rand_frames

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def O():
    from oracle import oracle
    return oracle


FORMATS = ["GRAY", "RGB", "RGBA", "NV21", "NV12", "YV12", "YV21"]
SIZES = [(200, 150), (130, 66), (1920, 1080), (97, 65)]



@pytest.mark.parametrize("angle", [90, 180, 270])
@pytest.mark.parametrize("fmt_name", FORMATS)
@pytest.mark.parametrize("W,H", SIZES)
def test_rotate_across_tiles(O, angle, fmt_name, W, H):
    fmt = getattr(O, "FMT_" + fmt_name)
    fr = rand_frames(O, 1, W, H, fmt, 1300 + W)[0]
    st, want, winfo = O.run_op((O.OP_ROTATE, angle), fr, W, H, fmt)
    st2, got, info = gpu_op((O.OP_ROTATE, angle), fr, W, H, fmt)
    assert st == st2
    if st == 0:
        assert info[:3] == tuple(winfo)
        assert_same_bits(got, want, "rotate %d %s %dx%d" % (angle, fmt_name, W, H))


@pytest.mark.parametrize("hv", [(1, 0), (0, 1), (1, 1)])
@pytest.mark.parametrize("fmt_name", FORMATS)
@pytest.mark.parametrize("W,H", SIZES[:3])
def test_flip_across_tiles(O, hv, fmt_name, W, H):
    fmt = getattr(O, "FMT_" + fmt_name)
    fr = rand_frames(O, 1, W, H, fmt, 1400 + W)[0]
    st, want, winfo = O.run_op((O.OP_FLIP,) + hv, fr, W, H, fmt)
    st2, got, info = gpu_op((O.OP_FLIP,) + hv, fr, W, H, fmt)
    assert st == st2
    if st == 0:
        assert info[:3] == tuple(winfo)
        assert_same_bits(got, want, "flip %s %s %dx%d" % (hv, fmt_name, W, H))


@pytest.mark.parametrize("box", [(3, 5, 170, 120), (64, 32, 191, 95), (1, 0, 198, 149), (0, 0, 199, 149), (33, 31, 100, 99)])
@pytest.mark.parametrize("fmt_name", FORMATS)
def test_crop_across_tiles(O, box, fmt_name):
    W, H = 200, 150
    fmt = getattr(O, "FMT_" + fmt_name)
    fr = rand_frames(O, 1, W, H, fmt, 1500)[0]
    st, want, winfo = O.run_op((O.OP_CROP,) + box, fr, W, H, fmt)
    st2, got, info = gpu_op((O.OP_CROP,) + box, fr, W, H, fmt)
    assert st == st2
    if st == 0:
        assert info[:3] == tuple(winfo)
        assert_same_bits(got, want, "crop %s %s" % (box, fmt_name))


@pytest.mark.parametrize("fmt_name", ["RGB", "NV21", "YV12"])
def test_crop_rotate_flip_chain(O, fmt_name):
    """Three index-map operators compose into one map: one tile-staged copy."""
    W, H = 320, 240
    fmt = getattr(O, "FMT_" + fmt_name)
    fr = rand_frames(O, 1, W, H, fmt, 1600)[0]
    ops = [(O.OP_CROP, 16, 8, 271, 199), (O.OP_ROTATE, 90), (O.OP_FLIP, 1, 0)]
    cur, w, h, f = fr, W, H, fmt
    for op in ops:
        st, cur, (w, h, f) = O.run_op(op, cur, w, h, f)
        assert st == 0
    st, got, info = gpu_image_chain(ops, fr, W, H, fmt, fmt, O.DT_U8, w, h)
    assert st == 0 and info[:3] == (w, h, fmt)
    assert_same_bits(got, cur, "crop+rotate+flip %s" % fmt_name)
