"""Camera-side layouts: Android `Image.Plane` row / pixel strides through
BandBufferSetFromYUVData / BandBufferSetFromDeviceYUVData (reference: Buffer::CreateFromYUVPlanes, buffer.cc:80-104;
JNI caller native_buffer_wrapper_jni.cc:34-67). The same pixels laid out with padded rows, with chroma in two
separate pixel-stride-2 buffers (YUV_420_888 as many devices deliver it) or as one interleaved buffer must give the
bits the oracle computes from the contiguous frame."""
import ctypes as C

import numpy as np
import pytest

from gpu_util import assert_same_bits, oracle_chain, to_capi_ops, NP_OF_ORC

pytestmark = pytest.mark.gpu
NORM = (5, 0, 0, 0, 0, 127.5, 127.5)


def _planes(rng, W, H):
    cw, ch = (W + 1) // 2, (H + 1) // 2
    return (rng.integers(0, 256, size=(H, W), dtype=np.uint8), rng.integers(0, 256, size=(ch, cw), dtype=np.uint8),
            rng.integers(0, 256, size=(ch, cw), dtype=np.uint8))


def _contiguous(O, fmt_name, Y, U, V):
    if fmt_name == "NV12":
        c = np.stack([U, V], axis=-1).reshape(U.shape[0], -1)
        return np.concatenate([Y.ravel(), c.ravel()])
    if fmt_name == "NV21":
        c = np.stack([V, U], axis=-1).reshape(U.shape[0], -1)
        return np.concatenate([Y.ravel(), c.ravel()])
    if fmt_name == "YV12":
        return np.concatenate([Y.ravel(), V.ravel(), U.ravel()])
    return np.concatenate([Y.ravel(), U.ravel(), V.ravel()])  # YV21


def _strided(Y, U, V, layout, y_pad, c_pad):
    """-> (backing arrays, y_off, u_off, v_off, row_stride_y, row_stride_uv, pixel_stride_uv); offsets into one
    backing array per plane group so device copies are simple."""
    H, W = Y.shape
    ch, cw = U.shape
    ys = W + y_pad
    ybuf = np.full((H, ys), 0xEE, dtype=np.uint8)
    ybuf[:, :W] = Y
    if layout == "planar":
        cs = cw + c_pad
        ub = np.full((ch, cs), 0xDD, dtype=np.uint8); ub[:, :cw] = U
        vb = np.full((ch, cs), 0xCC, dtype=np.uint8); vb[:, :cw] = V
        return ybuf, ub, vb, ys, cs, 1, 0
    cs = 2 * cw + c_pad
    if layout == "split_ps2":       # two independent buffers, samples every second byte, junk in between
        ub = np.full((ch, cs), 0xDD, dtype=np.uint8); ub[:, 0:2 * cw:2] = U
        vb = np.full((ch, cs), 0xCC, dtype=np.uint8); vb[:, 0:2 * cw:2] = V
        return ybuf, ub, vb, ys, cs, 2, 0
    # one interleaved buffer: U first ("uv") or V first ("vu"); the second pointer is the first + 1
    cbuf = np.full((ch, cs), 0xBB, dtype=np.uint8)
    first, second = (U, V) if layout == "uv" else (V, U)
    cbuf[:, 0:2 * cw:2] = first
    cbuf[:, 1:2 * cw:2] = second
    return ybuf, cbuf, None, ys, cs, 2, (1 if layout == "uv" else -1)


@pytest.mark.parametrize("fmt_name,layout,W,H,y_pad,c_pad", [
    ("NV21", "vu", 640, 480, 64, 64),          # padded rows, still 16-byte granular: row-staged kernel, bulk path
    ("NV12", "uv", 640, 480, 24, 8),           # word-granular strides: row-staged kernel, byte staging
    ("NV12", "split_ps2", 640, 480, 0, 16),    # YUV_420_888 with pixelStride 2 in two buffers: generic kernel
    ("YV12", "planar", 640, 480, 128, 64),     # padded planar
    ("YV21", "planar", 322, 242, 6, 3),        # odd strides and an odd-ish size
    ("NV21", "vu", 321, 241, 7, 5),            # odd width / height: chroma dims round up
])
@pytest.mark.parametrize("device", [False, True])
def test_strided_yuv_planes(O, fmt_name, layout, W, H, y_pad, c_pad, device):
    from band_b200 import band_c as B
    rng = np.random.default_rng(W * 7 + H)
    Y, U, V = _planes(rng, W, H)
    fmt = getattr(O, "FMT_" + fmt_name)
    frame = _contiguous(O, fmt_name, Y, U, V)
    assert frame.size == O.frame_bytes(W, H, fmt)
    ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, 160, 120), NORM]
    st, want = oracle_chain(O, ops, frame, W, H, fmt, O.DT_F32, (1, 120, 160, 3))
    assert st == 0
    ybuf, ub, vb, ys, cs, ps, second = _strided(Y, U, V, layout, y_pad, c_pad)
    L = B.lib()
    keep = [ybuf, ub, vb]
    frees = []

    def place(a):
        if a is None:
            return None
        a = np.ascontiguousarray(a)
        if not device:
            keep.append(a)
            return a.ctypes.data
        d = L.BandDeviceAlloc(a.nbytes)
        assert d and L.BandCopyToDevice(d, a.ctypes.data, a.nbytes) == B.kBandOk
        frees.append(d)
        return d
    yp, up = place(ybuf), place(ub)
    if vb is not None:
        vp = place(vb)
    elif second == 1:      # buffer holds U,V,U,V: u = base, v = base + 1
        vp = up + 1
    else:                  # buffer holds V,U,V,U: v = base, u = base + 1
        vp, up = up, up + 1
    try:
        buf = B.Buffer().set_yuv(yp, up, vp, W, H, ys, cs, ps, B.FORMATS[fmt_name], device=device)
        proc = B.ImageProcessor(to_capi_ops(ops))
        t = B.Tensor(np.float32, (1, 120, 160, 3), device=device)
        assert proc.process(buf, t) == B.kBandOk, B.last_error()
        assert_same_bits(t.numpy(), want[0], "%s %s device=%s" % (fmt_name, layout, device))
        # the same buffer through a chain with an index map (tile-staged or generic kernel)
        ops2 = [(O.OP_ROTATE, 90), (O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, 100, 150)]
        st, want2 = oracle_chain(O, ops2, frame, W, H, fmt, O.DT_U8, (1, 150, 100, 3))
        assert st == 0
        proc2 = B.ImageProcessor(to_capi_ops(ops2))
        t2 = B.Tensor(np.uint8, (1, 150, 100, 3), device=device)
        assert proc2.process(buf, t2) == B.kBandOk, B.last_error()
        assert_same_bits(t2.numpy(), want2[0], "rotated %s %s device=%s" % (fmt_name, layout, device))
        for x in (buf, t, t2, proc, proc2):
            x.close()
    finally:
        for d in frees:
            L.BandDeviceFree(d)
