"""Tile-staged kernel (band_b200/csrc/cuda/bandcu_tile.cu): 4:2:0 -> RGB -> bilinear down [-> Normalize] behind
Crop / Rotate / Flip index maps, planar and semi-planar chroma, both staging granules (16-byte cp.async and 32-bit
words), both lane mappings (quarter-turn maps walk output columns), partial edge tiles. Bit-compared with the oracle
and with the generic pixel-per-thread kernel (BANDCU_DISABLE_TILE=1) on the same inputs."""
import os

import numpy as np
import pytest

from gpu_util import assert_same_bits, gpu_chain, oracle_chain, rand_frames

pytestmark = pytest.mark.gpu
NORM = (5, 0, 0, 0, 0, 127.5, 127.5)


def _run_both(O, ops, fr, W, H, fmt, dt, dims, resident):
    st, want = oracle_chain(O, ops, fr, W, H, fmt, dt, dims)
    assert st == 0, st
    st, got = gpu_chain(ops, fr, W, H, fmt, dt, dims, device_resident=resident)
    assert st == 0
    assert_same_bits(got, want, "tile kernel vs oracle")
    os.environ["BANDCU_DISABLE_TILE"] = "1"   # read when the plan is created
    try:
        st, gen = gpu_chain(ops, fr, W, H, fmt, dt, dims, device_resident=resident)
    finally:
        del os.environ["BANDCU_DISABLE_TILE"]
    assert st == 0
    assert_same_bits(gen, want, "generic kernel vs oracle")


@pytest.mark.parametrize("fmt_name", ["YV12", "YV21", "NV12", "NV21"])
@pytest.mark.parametrize("angle", [0, 90, 180, 270])
@pytest.mark.parametrize("f32", [True, False])
def test_rotate_csc_resize_every_layout(O, fmt_name, angle, f32):
    """640x352 (16-byte granule): every 4:2:0 layout x every rotation, output 150x97 (partial edge tiles)."""
    W, H = 640, 352
    fmt = getattr(O, "FMT_" + fmt_name)
    ops = ([(O.OP_ROTATE, angle)] if angle else []) + [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, 150, 97)] + ([NORM] if f32 else [])
    fr = rand_frames(O, 3, W, H, fmt, 1000 + angle)
    _run_both(O, ops, fr, W, H, fmt, O.DT_F32 if f32 else O.DT_U8, (1, 97, 150, 3), resident=True)


@pytest.mark.parametrize("fmt_name,W,H", [("YV12", 648, 360), ("NV21", 324, 180), ("YV21", 328, 250)])
@pytest.mark.parametrize("angle", [90, 270])
def test_word_granule_and_odd_crops(O, fmt_name, W, H, angle):
    """Widths that are multiples of 4/8 but not of 16/32 take the 32-bit staging path; the crop starts on odd
    luma coordinates, so the chroma sample of a pixel is not the one under it in the source."""
    fmt = getattr(O, "FMT_" + fmt_name)
    # after a quarter-turn the frame is H wide and W high
    ops = [(O.OP_ROTATE, angle), (O.OP_CROP, 3, 5, H - 8, W - 12), (O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, 96, 130), NORM]
    fr = rand_frames(O, 2, W, H, fmt, 77)
    _run_both(O, ops, fr, W, H, fmt, O.DT_F32, (1, 130, 96, 3), resident=False)


@pytest.mark.parametrize("fmt_name", ["YV12", "NV12"])
@pytest.mark.parametrize("flip", [(1, 0), (0, 1), (1, 1)])
def test_flip_crop_csc_resize(O, fmt_name, flip):
    W, H = 1280, 720
    fmt = getattr(O, "FMT_" + fmt_name)
    ops = [(O.OP_FLIP, flip[0], flip[1]), (O.OP_CROP, 100, 50, 1179, 669), (O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, 300, 300)]
    fr = rand_frames(O, 2, W, H, fmt, 31)
    _run_both(O, ops, fr, W, H, fmt, O.DT_U8, (1, 300, 300, 3), resident=True)


def test_planar_1080p_to_224_f32_identity_map(O):
    """Planar 4:2:0 without any index map (the row-staged kernel only takes semi-planar chroma)."""
    ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, 224, 224), NORM]
    fr = rand_frames(O, 3, 1920, 1080, O.FMT_YV12, 8)
    _run_both(O, ops, fr, 1920, 1080, O.FMT_YV12, O.DT_F32, (1, 224, 224, 3), resident=True)


def test_saturating_frames_through_the_tile_kernel(O):
    """All-0 / all-255 / extreme-chroma frames exercise every clamp of the packed YuvPixel arithmetic."""
    W, H = 640, 480
    fb = O.frame_bytes(W, H, O.FMT_YV12)
    fr = np.zeros((4, fb), dtype=np.uint8)
    fr[1] = 255
    fr[2, :W * H] = 255            # Y = 255, U = V = 0
    fr[3, W * H:] = 255            # Y = 0,   U = V = 255
    ops = [(O.OP_ROTATE, 90), (O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, 100, 120)]
    _run_both(O, ops, fr, W, H, O.FMT_YV12, O.DT_U8, (1, 120, 100, 3), resident=False)


@pytest.mark.parametrize("fmt_name,W,H,dw,dh,f32", [
    ("YV12", 1920, 1080, 300, 300, False),   # planar chroma through the ROW-staged kernel (two bulk copies per chroma row)
    ("YV21", 1280, 720, 513, 400, True),
    ("YV12", 648, 360, 111, 97, True),       # 324-byte chroma rows: not 16-byte granular -> producer-warp byte staging
    ("YV21", 642, 362, 100, 90, False),      # width not a multiple of 4: neither staged kernel's bulk path applies
])
def test_planar_identity_map_row_staged(O, fmt_name, W, H, dw, dh, f32):
    fmt = getattr(O, "FMT_" + fmt_name)
    ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, dw, dh)] + ([NORM] if f32 else [])
    fr = rand_frames(O, 5, W, H, fmt, 4321)
    dt = O.DT_F32 if f32 else O.DT_U8
    st, want = oracle_chain(O, ops, fr, W, H, fmt, dt, (1, dh, dw, 3))
    assert st == 0
    for resident in (True, False):
        st, got = gpu_chain(ops, fr, W, H, fmt, dt, (1, dh, dw, 3), device_resident=resident)
        assert st == 0
        assert_same_bits(got, want, "planar %s resident=%s" % (fmt_name, resident))
    os.environ["BANDCU_DISABLE_FAST"] = "1"
    try:
        st, got = gpu_chain(ops, fr, W, H, fmt, dt, (1, dh, dw, 3), device_resident=True)
    finally:
        del os.environ["BANDCU_DISABLE_FAST"]
    assert st == 0
    assert_same_bits(got, want, "planar %s without the row-staged kernel" % fmt_name)


@pytest.mark.parametrize("fmt_name", ["YV12", "NV21"])
@pytest.mark.parametrize("pre", [[(2, 90)], [(3, 1, 0)], [(2, 270), (3, 0, 1)]])
def test_mild_down_scale_stages_whole_rectangles(O, fmt_name, pre):
    """Ratios below 2 keep rectangle staging of the luma plane (consecutive samples share source rows); ratios of 2
    and more stage per sample. 640x352 -> 250x400 after a quarter-turn is 1.41 x 1.6; the flip case is 1.6 x 1.41 on
    one axis and crosses the threshold on neither."""
    W, H = 640, 352
    fmt = getattr(O, "FMT_" + fmt_name)
    turned = any(op[0] == 2 and op[1] in (90, 270) for op in pre)
    dw, dh = (250, 400) if turned else (400, 250)
    ops = list(pre) + [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, dw, dh), NORM]
    fr = rand_frames(O, 2, W, H, fmt, 99)
    _run_both(O, ops, fr, W, H, fmt, O.DT_F32, (1, dh, dw, 3), resident=True)
    # and the same chain with per-sample staging switched off must agree too (both modes, same bits)
    os.environ["BANDCU_TILE_NO_SKIP"] = "1"
    try:
        ops2 = list(pre) + [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, dw // 3, dh // 3)]
        st, want = oracle_chain(O, ops2, fr, W, H, fmt, O.DT_U8, (1, dh // 3, dw // 3, 3))
        st2, got = gpu_chain(ops2, fr, W, H, fmt, O.DT_U8, (1, dh // 3, dw // 3, 3), device_resident=True)
    finally:
        del os.environ["BANDCU_TILE_NO_SKIP"]
    assert st == 0 and st2 == 0
    assert_same_bits(got, want, "rectangle staging at a large ratio")


@pytest.mark.parametrize("fmt_name", ["NV12", "YV12"])
@pytest.mark.parametrize("order,box", [
    ("crop_first", (64, 32, 1343, 751)),    # 16-byte aligned origin: row-staged kernel, bulk path (center-crop style)
    ("crop_first", (3, 5, 1202, 700)),      # odd origin: chroma origin is (x0/2, y0/2); byte staging
    ("csc_first", (128, 64, 1151, 655)),    # crop AFTER the conversion, even origin: same samples as a chroma translation
    ("csc_first", (5, 3, 1004, 602)),       # odd origin after the conversion: chroma phase differs -> tile-staged kernel
])
def test_crop_folds_into_the_row_staged_kernel(O, fmt_name, order, box):
    W, H = 1408, 800
    fmt = getattr(O, "FMT_" + fmt_name)
    crop = (O.OP_CROP,) + box
    csc = (O.OP_CSC, O.FMT_RGB)
    ops = ([crop, csc] if order == "crop_first" else [csc, crop]) + [(O.OP_RESIZE, 224, 224), NORM]
    fr = rand_frames(O, 3, W, H, fmt, 555)
    st, want = oracle_chain(O, ops, fr, W, H, fmt, O.DT_F32, (1, 224, 224, 3))
    assert st == 0
    for resident in (True, False):
        st, got = gpu_chain(ops, fr, W, H, fmt, O.DT_F32, (1, 224, 224, 3), device_resident=resident)
        assert st == 0
        assert_same_bits(got, want, "%s %s resident=%s" % (order, box, resident))
    os.environ["BANDCU_DISABLE_FAST"] = "1"
    try:
        st, got = gpu_chain(ops, fr, W, H, fmt, O.DT_F32, (1, 224, 224, 3), device_resident=True)
    finally:
        del os.environ["BANDCU_DISABLE_FAST"]
    assert st == 0
    assert_same_bits(got, want, "%s %s without the row-staged kernel" % (order, box))


@pytest.mark.parametrize("fmt_name", ["NV12", "YV21"])
@pytest.mark.parametrize("pre,dw,dh", [([(2, 90)], 50, 100), ([(3, 1, 1)], 120, 44), ([(2, 270), (0, 8, 16, 327, 591)], 40, 64)])
def test_large_ratios_stage_chroma_per_sample_too(O, fmt_name, pre, dw, dh):
    """Ratios of 6 and more along the row-mapped axis: luma AND chroma planes are staged per sample."""
    W, H = 640, 352
    fmt = getattr(O, "FMT_" + fmt_name)
    ops = list(pre) + [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, dw, dh), NORM]
    fr = rand_frames(O, 2, W, H, fmt, 606)
    _run_both(O, ops, fr, W, H, fmt, O.DT_F32, (1, dh, dw, 3), resident=True)
