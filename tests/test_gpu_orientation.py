"""AutoOrient (ADDITIVE): the EXIF orientation a Buffer carries (band/common.h:146-161, never applied by
the reference) applied as part of the fused index map. Checked (1) against numpy's definition of the eight EXIF
transforms on RGB images, (2) on 4:2:0 frames against the oracle running the equivalent Flip / Rotate operators."""
import numpy as np
import pytest

from gpu_util import assert_same_bits, oracle_chain, to_capi_ops

pytestmark = pytest.mark.gpu
NORM = (5, 0, 0, 0, 0, 127.5, 127.5)


def _upright(a, o):
    """EXIF orientation o of stored array a (H, W, C) -> the upright image."""
    return {1: a, 2: a[:, ::-1], 3: a[::-1, ::-1], 4: a[::-1], 5: a.transpose(1, 0, 2), 6: np.rot90(a, -1),
            7: a[::-1, ::-1].transpose(1, 0, 2), 8: np.rot90(a, 1)}[o]


@pytest.mark.parametrize("o", range(1, 9))
@pytest.mark.parametrize("device", [False, True])
def test_auto_orient_rgb_matches_exif_definition(o, device):
    from band_b200 import band_c as B
    W, H = 52, 37
    rng = np.random.default_rng(o)
    img = rng.integers(0, 256, size=(H, W, 3), dtype=np.uint8)
    want = np.ascontiguousarray(_upright(img, o))
    L = B.lib()
    dev = None
    if device:
        dev = L.BandDeviceAlloc(img.nbytes)
        assert dev and L.BandCopyToDevice(dev, img.ctypes.data, img.nbytes) == B.kBandOk
        buf = B.Buffer().set_raw(dev, W, H, B.kBandRGB, device=True)
    else:
        buf = B.Buffer.from_numpy(img.reshape(-1), W, H, B.kBandRGB)
    buf.set_orientation(o)
    proc = B.ImageProcessor([(B.BAND_AUTO_ORIENT,)])
    t = B.Tensor(np.uint8, (1,) + want.shape, device=device)
    assert proc.process(buf, t) == B.kBandOk, B.last_error()
    assert_same_bits(t.numpy()[0], want, "orientation %d" % o)
    # the same processor re-plans when the next frame carries another orientation
    o2 = 1 + (o % 8)
    if _upright(img, o2).shape == want.shape:
        buf.set_orientation(o2)
        assert proc.process(buf, t) == B.kBandOk, B.last_error()
        assert_same_bits(t.numpy()[0], np.ascontiguousarray(_upright(img, o2)), "orientation %d after %d" % (o2, o))
    for x in (buf, t, proc):
        x.close()
    if dev:
        L.BandDeviceFree(dev)


@pytest.mark.parametrize("fmt_name", ["NV21", "YV12"])
@pytest.mark.parametrize("o", range(1, 9))
def test_auto_orient_yuv_chain_equals_flip_rotate(O, fmt_name, o):
    from band_b200 import band_c as B
    W, H = 640, 480
    fmt = getattr(O, "FMT_" + fmt_name)
    rng = np.random.default_rng(100 + o)
    frame = rng.integers(0, 256, size=O.frame_bytes(W, H, fmt), dtype=np.uint8)
    mirror, angle = {1: (0, 0), 2: (1, 0), 3: (0, 180), 4: (1, 180), 5: (1, 90), 6: (0, 270), 7: (1, 270), 8: (0, 90)}[o]
    pre = ([(O.OP_FLIP, 1, 0)] if mirror else []) + ([(O.OP_ROTATE, angle)] if angle else [])
    tail = [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, 200, 160), NORM]
    st, want = oracle_chain(O, pre + tail, frame, W, H, fmt, O.DT_F32, (1, 160, 200, 3))
    assert st == 0
    buf = B.Buffer.from_numpy(frame, W, H, B.FORMATS[fmt_name]).set_orientation(o)
    proc = B.ImageProcessor([(B.BAND_AUTO_ORIENT,)] + to_capi_ops(tail))
    t = B.Tensor(np.float32, (1, 160, 200, 3), device=True)
    assert proc.process(buf, t) == B.kBandOk, B.last_error()
    assert_same_bits(t.numpy(), want[0], "%s orientation %d" % (fmt_name, o))
    for x in (buf, t, proc):
        x.close()


def test_orientation_argument_checks():
    from band_b200 import band_c as B
    L = B.lib()
    img = np.zeros((4, 4, 3), dtype=np.uint8)
    buf = B.Buffer.from_numpy(img.reshape(-1), 4, 4, B.kBandRGB)
    assert L.BandBufferSetOrientation(buf.handle, 0) == B.kBandErr
    assert L.BandBufferSetOrientation(buf.handle, 9) == B.kBandErr
    assert L.BandBufferSetOrientation(None, 1) == B.kBandErr
    empty = B.Buffer()
    assert L.BandBufferSetOrientation(empty.handle, 1) == B.kBandErr   # no data set yet
    # a tensor of the un-turned shape is rejected for a quarter-turn orientation
    buf2 = B.Buffer.from_numpy(np.zeros(6 * 4 * 3, dtype=np.uint8), 6, 4, B.kBandRGB).set_orientation(6)
    proc = B.ImageProcessor([(B.BAND_AUTO_ORIENT,)])
    t = B.Tensor(np.uint8, (1, 4, 6, 3), device=True)
    assert proc.process(buf2, t) == B.kBandErr
