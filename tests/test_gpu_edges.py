"""GPU parity: edge cases the domain has (odd / tiny / ragged dims, saturating inputs, every YUV triple,
dtype pairs of Normalize, status codes) and size-independent properties at BASELINE's full sizes."""
import ctypes as C

import numpy as np
import pytest

from gpu_util import assert_same_bits, gpu_chain, gpu_image_chain, gpu_op, oracle_chain, rand_frames

pytestmark = pytest.mark.gpu
NORM = (5, 0, 0, 0, 0, 127.5, 127.5)


@pytest.mark.parametrize("fmt_name", ["NV12", "NV21", "YV12", "YV21"])
@pytest.mark.parametrize("W,H,dw,dh", [(641, 479, 224, 224), (3, 5, 2, 3), (2, 2, 3, 2), (17, 9, 5, 4), (642, 482, 225, 223),
                                       (1921, 1081, 300, 300)])
def test_odd_and_tiny_yuv_frames(O, fmt_name, W, H, dw, dh):
    fmt = getattr(O, "FMT_" + fmt_name)
    ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, dw, dh), NORM]
    fr = rand_frames(O, 2, W, H, fmt, W * 7 + H)
    st, want = oracle_chain(O, ops, fr, W, H, fmt, O.DT_F32, (1, dh, dw, 3))
    assert st == 0
    st, got = gpu_chain(ops, fr, W, H, fmt, O.DT_F32, (1, dh, dw, 3))
    assert st == 0
    assert_same_bits(got, want, "%s %dx%d" % (fmt_name, W, H))


@pytest.mark.parametrize("fill", [0, 255, "checker", "ramp"])
def test_saturating_frames(O, fill):
    W, H = 1920, 1080
    fb = O.frame_bytes(W, H, O.FMT_NV12)
    if fill == "checker":
        fr = ((np.arange(fb) + (np.arange(fb) // W)) % 2 * 255).astype(np.uint8)
    elif fill == "ramp":
        fr = (np.arange(fb) % 256).astype(np.uint8)
    else:
        fr = np.full(fb, fill, dtype=np.uint8)
    ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, 224, 224), NORM]
    st, want = oracle_chain(O, ops, fr, W, H, O.FMT_NV12, O.DT_F32, (1, 224, 224, 3))
    st2, got = gpu_chain(ops, fr, W, H, O.FMT_NV12, O.DT_F32, (1, 224, 224, 3))
    assert st == 0 and st2 == 0
    assert_same_bits(got, want, str(fill))


def test_every_yuv_triple(O):
    """All 2^24 (Y,U,V) triples once: a 4096x4096 frame whose 2x2 blocks share one (U,V) and sweep Y
    through ColorSpaceConvert(kRGB) alone (planar YV21 so U and V are free planes)."""
    W = H = 8192
    yy = np.arange(256, dtype=np.uint8)
    # chroma plane 4096x4096: block (i,j) -> U = i % 256, V = j % 256 ; luma cycles so every Y meets every (U,V)
    u = (np.arange(4096) % 256).astype(np.uint8)[:, None].repeat(4096, 1)
    v = (np.arange(4096) % 256).astype(np.uint8)[None, :].repeat(4096, 0)
    ci = (np.arange(4096) // 256)[:, None] * 16 + (np.arange(4096) // 256)[None, :]  # 0..255: the Y of this block
    y = yy[ci].repeat(2, 0).repeat(2, 1)
    frame = np.concatenate([y.reshape(-1), u.reshape(-1), v.reshape(-1)])
    ops = [(O.OP_CSC, O.FMT_RGB)]
    st, want, _ = O.run_op(ops[0], frame, W, H, O.FMT_YV21)
    assert st == 0
    st, got, info = gpu_op(ops[0], frame, W, H, O.FMT_YV21, cap=W * H * 3)
    assert st == 0 and info[:3] == (W, H, O.FMT_RGB)
    assert_same_bits(got, want, "all YUV triples")


@pytest.mark.parametrize("out_dt", [1, 2, 3, 7, 9])
def test_normalize_dtype_pairs(O, out_dt):
    """common_operator.cc:42-65: u8 RGB -> every supported output element type (values kept in range so the
    cast of the difference is defined for unsigned outputs)."""
    W, H = 37, 23
    rng = np.random.default_rng(out_dt)
    fr = rng.integers(100, 200, size=W * H * 3, dtype=np.uint8)
    ops = [(O.OP_NORMALIZE, 0, 0, 0, 0, 64.0, 2.0)]
    st, want = oracle_chain(O, ops, fr, W, H, O.FMT_RGB, out_dt, (1, H, W, 3))
    st2, got = gpu_chain(ops, fr, W, H, O.FMT_RGB, out_dt, (1, H, W, 3))
    assert st == 0 and st2 == 0
    assert_same_bits(got, want, "normalize -> %d" % out_dt)


def test_data_type_convert_u8_to_f32(O):
    ops = [(O.OP_DTC,)]
    fr = rand_frames(O, 1, 64, 48, O.FMT_RGB, 9)
    st, want = oracle_chain(O, ops, fr, 64, 48, O.FMT_RGB, O.DT_F32, (1, 48, 64, 3))
    st2, got = gpu_chain(ops, fr, 64, 48, O.FMT_RGB, O.DT_F32, (1, 48, 64, 3))
    assert st == 0 and st2 == 0
    assert_same_bits(got, want, "DataTypeConvert")


@pytest.mark.parametrize("ops", [
    [(1, 64, 48)],                                   # Resize to identical dims (image_operator.cc:180-184)
    [(5, 0, 0, 0, 0, 127.5, 127.5)],                 # Normalize on YUV input (common_operator.cc:111-117)
    [(7,)],                                          # AutoConvert rejects YUV (image_operator.cc:493-501)
    [(0, 0, 0, 64, 10)],                             # Crop x1 == W out of bounds
])
def test_error_statuses_match_oracle(O, ops):
    fr = rand_frames(O, 1, 64, 48, O.FMT_NV12, 3)
    fmt = O.FMT_NV12
    odims = (1, 48, 64, 3)
    st, _ = O.run_chain(ops, fr[0], 64, 48, fmt, O.DT_U8, odims)
    assert st != 0
    st2, out = gpu_chain(ops, fr, 64, 48, fmt, O.DT_U8, odims)
    assert st2 != 0 and out is None


@pytest.mark.parametrize("angle", [0, 90, 180, 270, -90, 450])
@pytest.mark.parametrize("fmt_name,W,H", [("GRAY", 33, 21), ("RGB", 33, 21), ("RGBA", 20, 31), ("YV12", 35, 21), ("NV12", 34, 22),
                                          ("NV21", 33, 23)])
def test_rotate_all_formats(O, angle, fmt_name, W, H):
    fmt = getattr(O, "FMT_" + fmt_name)
    fr = rand_frames(O, 1, W, H, fmt, 77)[0]
    st, want, winfo = O.run_op((O.OP_ROTATE, angle), fr, W, H, fmt)
    st2, got, info = gpu_op((O.OP_ROTATE, angle), fr, W, H, fmt)
    assert st == st2
    if st == 0:
        assert info[:3] == tuple(winfo)
        assert_same_bits(got, want, "rotate %d %s" % (angle, fmt_name))


@pytest.mark.parametrize("fmt_name,W,H,box", [("GRAY", 33, 21, (3, 2, 20, 11)), ("RGB", 33, 21, (1, 1, 31, 19)),
                                              ("RGBA", 20, 31, (5, 7, 9, 30)), ("YV12", 36, 22, (3, 5, 30, 20)),
                                              ("NV12", 36, 22, (4, 2, 33, 19)), ("NV21", 37, 23, (1, 1, 36, 22))])
def test_crop_all_formats(O, fmt_name, W, H, box):
    fmt = getattr(O, "FMT_" + fmt_name)
    fr = rand_frames(O, 1, W, H, fmt, 78)[0]
    st, want, winfo = O.run_op((O.OP_CROP,) + box, fr, W, H, fmt)
    st2, got, info = gpu_op((O.OP_CROP,) + box, fr, W, H, fmt)
    assert st == 0 and st2 == 0 and info[:3] == tuple(winfo)
    assert_same_bits(got, want, "crop %s" % fmt_name)


@pytest.mark.parametrize("hv", [(1, 0), (0, 1), (1, 1)])
@pytest.mark.parametrize("fmt_name,W,H", [("GRAY", 33, 21), ("RGB", 33, 21), ("RGBA", 20, 31), ("YV12", 36, 22), ("NV12", 36, 22)])
def test_flip_all_formats(O, hv, fmt_name, W, H):
    fmt = getattr(O, "FMT_" + fmt_name)
    fr = rand_frames(O, 1, W, H, fmt, 79)[0]
    st, want, winfo = O.run_op((O.OP_FLIP,) + hv, fr, W, H, fmt)
    st2, got, info = gpu_op((O.OP_FLIP,) + hv, fr, W, H, fmt)
    assert st == st2
    if st == 0:
        assert info[:3] == tuple(winfo)
        assert_same_bits(got, want, "flip %s %s" % (hv, fmt_name))


@pytest.mark.parametrize("fmt_name", ["NV12", "NV21", "YV12", "YV21"])
@pytest.mark.parametrize("W,H,dw,dh", [(64, 48, 22, 18), (64, 48, 32, 24), (64, 48, 48, 36), (64, 48, 24, 18),
                                       (32, 24, 64, 48), (33, 25, 66, 50), (32, 24, 64, 24), (32, 24, 80, 70),
                                       (65, 47, 23, 21), (64, 48, 64, 20), (64, 48, 16, 12), (66, 54, 22, 18),
                                       (1920, 1080, 640, 360), (640, 480, 300, 300)])
def test_resize_in_yuv_domain(O, fmt_name, W, H, dw, dh):
    """ResizeYv / ResizeNv (libyuv_image_operator.cc:376-435): I420Scale = ScalePlane per plane, NV12Scale =
    ScalePlane(Y) + UVScale(interleaved chroma) incl. their Down2/Down34/Down38/Up2/vertical special cases."""
    fmt = getattr(O, "FMT_" + fmt_name)
    fr = rand_frames(O, 1, W, H, fmt, W + 3 * dw)[0]
    st, want, winfo = O.run_op((O.OP_RESIZE, dw, dh), fr, W, H, fmt)
    st2, got, info = gpu_op((O.OP_RESIZE, dw, dh), fr, W, H, fmt, cap=max(1 << 16, dw * dh * 4))
    assert st == 0 and st2 == 0, (st, st2)
    assert info[:3] == tuple(winfo)
    assert_same_bits(got, want, "yuv resize %s %dx%d->%dx%d" % (fmt_name, W, H, dw, dh))


@pytest.mark.parametrize("fmt_name", ["NV12", "NV21", "YV12", "YV21", "GRAY"])
@pytest.mark.parametrize("W,H,dw,dh", [
    (1920, 1080, 1280, 720), (1920, 1080, 960, 540), (1920, 1080, 1056, 594), (1920, 1080, 416, 234), (1920, 1080, 480, 270),
    (1282, 722, 801, 403),    # chroma rows of 641 bytes: partial last word, odd strides -> the fallback kernels
    (1284, 724, 803, 405),    # word-aligned strides, odd output width: byte stores at the row ends
    (648, 488, 640, 240),     # unit-ish horizontal step, 1/2 vertical
    (640, 480, 700, 300),     # up horizontally, down vertically: still Scale*BilinearDown
    (3840, 2160, 1280, 1080), (640, 480, 160, 120), (640, 480, 80, 120)])
def test_plane_down_scales_row_staged(O, fmt_name, W, H, dw, dh):
    """bandcu_planes.cu: one warp per output row, blended row pair in shared memory, then the column filter -- against the
    oracle's ScalePlane / UVScale (and, through BANDCU_DISABLE_PLANE_ROWS in the variants run, the pixel-per-thread kernel)."""
    fmt = getattr(O, "FMT_" + fmt_name)
    fr = rand_frames(O, 1, W, H, fmt, W + 5 * dw)[0]
    st, want, winfo = O.run_op((O.OP_RESIZE, dw, dh), fr, W, H, fmt)
    st2, got, info = gpu_op((O.OP_RESIZE, dw, dh), fr, W, H, fmt, cap=max(1 << 16, dw * dh * 4))
    assert st == 0 and st2 == 0, (st, st2)
    assert info[:3] == tuple(winfo)
    assert_same_bits(got, want, "plane rows %s %dx%d->%dx%d" % (fmt_name, W, H, dw, dh))


@pytest.mark.parametrize("fmt_name", ["NV12", "NV21", "YV12", "YV21", "GRAY"])
@pytest.mark.parametrize("W,H,dw,dh", [
    (960, 540, 1920, 1080), (642, 362, 1284, 724), (641, 361, 1282, 722), (641, 361, 1281, 721), (642, 362, 1283, 723),
    (100, 60, 200, 60), (100, 60, 199, 60), (102, 62, 204, 186),   # rows kept / tripled: the Linear form
    (6, 4, 12, 8), (2, 2, 4, 4), (3, 3, 5, 5), (1920, 32, 3840, 64)])
def test_plane_exact_2x_up_samplers(O, fmt_name, W, H, dw, dh):
    """ScalePlaneUp2_{Linear,Bilinear} / ScaleUV{Linear,Bilinear}Up2 and their edge rules (bandcu_planes.cu plane_up2_kernel)."""
    fmt = getattr(O, "FMT_" + fmt_name)
    fr = rand_frames(O, 1, W, H, fmt, W + 7 * dw)[0]
    st, want, winfo = O.run_op((O.OP_RESIZE, dw, dh), fr, W, H, fmt)
    st2, got, info = gpu_op((O.OP_RESIZE, dw, dh), fr, W, H, fmt, cap=max(1 << 16, dw * dh * 4))
    assert st == 0 and st2 == 0, (st, st2)
    assert info[:3] == tuple(winfo)
    assert_same_bits(got, want, "up2 %s %dx%d->%dx%d" % (fmt_name, W, H, dw, dh))


@pytest.mark.parametrize("src_name", ["RGBA", "RGB"])
@pytest.mark.parametrize("W,H,dw,dh", [(64, 64, 128, 128), (100, 60, 200, 120), (33, 21, 65, 41), (33, 21, 66, 42), (50, 40, 100, 40)])
def test_gray_conversion_then_exact_2x_up(O, src_name, W, H, dw, dh):
    """CSC(Gray) -> Resize at exactly 2x: RAWToJ400 / ARGBToJ400 on the fetch, then ScalePlaneUp2 (BASELINE configs[4]'s
    RGBA 64^2 -> gray 128^2 point)."""
    fmt = getattr(O, "FMT_" + src_name)
    ops = [(O.OP_CSC, O.FMT_GRAY), (O.OP_RESIZE, dw, dh)]
    fr = rand_frames(O, 3, W, H, fmt, 17 + W)
    cur_w = []
    for f in fr:
        cur, w, h, ff = f, W, H, fmt
        for op in ops:
            st, cur, (w, h, ff) = O.run_op(op, cur, w, h, ff)
            assert st == 0
        cur_w.append(cur)
    for k, f in enumerate(fr):
        st, got, info = gpu_image_chain(ops, f, W, H, fmt, O.FMT_GRAY, O.DT_U8, dw, dh)
        assert st == 0 and info[:3] == (dw, dh, O.FMT_GRAY)
        assert_same_bits(got, cur_w[k], "%s -> gray -> %dx%d" % (src_name, dw, dh))


@pytest.mark.parametrize("fmt_name", ["NV21", "YV12"])
def test_resize_yuv_then_convert_chain(O, fmt_name):
    """Chain order matters: Resize in the YUV domain THEN ColorSpaceConvert uses the plane blenders, not the ARGB one."""
    fmt = getattr(O, "FMT_" + fmt_name)
    W, H, d = 640, 480, 224
    ops = [(O.OP_RESIZE, d, d), (O.OP_CSC, O.FMT_RGB), NORM]
    fr = rand_frames(O, 2, W, H, fmt, 31)
    st, want = oracle_chain(O, ops, fr, W, H, fmt, O.DT_F32, (1, d, d, 3))
    st2, got = gpu_chain(ops, fr, W, H, fmt, O.DT_F32, (1, d, d, 3))
    assert st == 0 and st2 == 0
    assert_same_bits(got, want, "resize(yuv)->csc " + fmt_name)


@pytest.mark.parametrize("dst_name", ["NV12", "NV21", "YV12", "YV21"])
@pytest.mark.parametrize("src_name,W,H", [("RGB", 64, 48), ("RGBA", 64, 48), ("RGB", 65, 47), ("RGBA", 33, 21), ("RGB", 2, 2),
                                          ("RGBA", 1, 1), ("RGB", 1920, 1080)])
def test_rgb_to_yuv_conversions(O, src_name, dst_name, W, H):
    """ConvertFromRgb / ConvertFromRgba -> 4:2:0 (libyuv_image_operator.cc:523-587,621-674): RAWToI420, ABGRToNV12,
    ABGRToNV21, ABGRToI420 with libyuv's default C rows (parity unpinned: no reference test covers them)."""
    sf, df = getattr(O, "FMT_" + src_name), getattr(O, "FMT_" + dst_name)
    fr = rand_frames(O, 1, W, H, sf, 5 * W + H)[0]
    st, want, winfo = O.run_op((O.OP_CSC, df), fr, W, H, sf)
    st2, got, info = gpu_op((O.OP_CSC, df), fr, W, H, sf)
    assert st == 0 and st2 == 0, (st, st2)
    assert info[:3] == tuple(winfo)
    assert_same_bits(got, want, "%s->%s %dx%d" % (src_name, dst_name, W, H))


@pytest.mark.parametrize("src_name", ["NV12", "NV21", "YV12", "YV21"])
@pytest.mark.parametrize("dst_name", ["NV12", "NV21", "YV12", "YV21", "RGBA", "GRAY"])
def test_yuv_repacks_and_other_targets(O, src_name, dst_name):
    """YUV <-> YUV repacks (:125-162,:213-250,:302-367), YUV -> RGBA (..ToABGR) and YUV -> gray (Y-plane copy)."""
    if src_name == dst_name:
        pytest.skip("same format is an error in both implementations (checked elsewhere)")
    sf, df = getattr(O, "FMT_" + src_name), getattr(O, "FMT_" + dst_name)
    W, H = 37, 23
    fr = rand_frames(O, 1, W, H, sf, 99)[0]
    st, want, winfo = O.run_op((O.OP_CSC, df), fr, W, H, sf)
    st2, got, info = gpu_op((O.OP_CSC, df), fr, W, H, sf)
    assert st == st2
    if st == 0:
        assert info[:3] == tuple(winfo)
        assert_same_bits(got, want, "%s->%s" % (src_name, dst_name))


def test_crop_then_rgb_to_yuv_chain(O):
    """A pending index map on the packed source folds into the RGB -> YUV launch."""
    W, H = 70, 50
    fr = rand_frames(O, 1, W, H, O.FMT_RGB, 5)[0]
    ops = [(O.OP_ROTATE, 90), (O.OP_CROP, 3, 5, 40, 60), (O.OP_CSC, O.FMT_NV21)]
    cur, w, h, f = fr, W, H, O.FMT_RGB
    for op in ops:
        st, cur, (w, h, f) = O.run_op(op, cur, w, h, f)
        assert st == 0
    st, got, info = gpu_image_chain(ops, fr, W, H, O.FMT_RGB, O.FMT_NV21, O.DT_U8, w, h)
    assert st == 0 and info[:3] == (w, h, O.FMT_NV21)
    assert_same_bits(got, cur, "rotate+crop+rgb->nv21")


def test_cpp_benchmark_tool_runs_the_fused_path(tmp_path):
    """tools/band_preproc_benchmark.cc: the C++ CLI (a client of the two C ABIs only) runs a small config on the GPU and
    reports one fused launch per step."""
    import json
    import os
    import subprocess
    from conftest import ROOT
    exe = os.path.join(ROOT, "band_b200", "lib", "band_preproc_benchmark")
    assert os.path.exists(exe), "run __graft_entry__.build()"
    cfg = tmp_path / "cfg.json"
    cfg.write_text(json.dumps({"format": "NV21", "width": 640, "height": 480,
                               "operators": [["color_space_convert", "RGB"], ["resize", 224, 224], ["normalize", 127.5, 127.5]],
                               "output": {"dtype": "float32", "dims": [1, 224, 224, 3]},
                               "frames_per_step": 64, "steps": 3, "warmup": 1, "devices": [0]}))
    out = subprocess.run([exe, str(cfg)], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    rep = json.loads(out.stdout.strip().splitlines()[-1])
    assert rep["fused_single_launch"] == 1 and rep["launches_per_step"] == 1.0 and rep["frames_per_s"] > 0
    bad = tmp_path / "bad.json"
    bad.write_text('{"operators": [["sharpen", 1]]}')
    out = subprocess.run([exe, str(bad)], capture_output=True, text=True, timeout=60)
    assert out.returncode == 2 and "unknown operator" in out.stderr


def test_recorded_batch_replays_with_new_frame_contents(O):
    """BandImageProcessorRecordBatch: one frame recorded into a CUDA graph; the replay sees whatever the frame memory holds
    at replay time (a camera ring slot refilled in place) and produces the oracle's bits. Pointer-table batches and
    pageable host frames are refused."""
    from band_b200 import band_c as B
    from gpu_util import rand_frames, oracle_chain, to_capi_ops, assert_same_bits
    W, H, d = 640, 480, 224
    ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, d, d), (O.OP_NORMALIZE, 0, 0, 0, 0, 127.5, 127.5)]
    fr = rand_frames(O, 3, W, H, O.FMT_NV21, 2024)
    st, want = oracle_chain(O, ops, fr, W, H, O.FMT_NV21, O.DT_F32, (1, d, d, 3))
    assert st == 0
    L = B.lib()
    fb = fr.shape[1]
    dev = L.BandDeviceAlloc(fb)
    assert dev and L.BandCopyToDevice(dev, fr[0].ctypes.data, fb) == B.kBandOk
    buf = B.Buffer().set_raw(dev, W, H, B.kBandNV21, device=True)
    t = B.Tensor(np.float32, (1, d, d, 3), device=True)
    proc = B.ImageProcessor(to_capi_ops(ops))
    rec = proc.record_batch([buf], [t])
    try:
        assert_same_bits(t.numpy(), want[0], "the recording pass itself")
        for k in (1, 2, 0):
            assert L.BandCopyToDevice(dev, fr[k].ctypes.data, fb) == B.kBandOk
            n0 = L.BandCudaLaunchCount()
            assert proc.replay(rec) == B.kBandOk, B.last_error()
            assert L.BandCudaLaunchCount() - n0 == 1
            assert_same_bits(t.numpy(), want[k], "replay %d" % k)
        # ordinary calls still work next to a recording
        assert proc.process(buf, t) == B.kBandOk
    finally:
        proc.release(rec)
    # refused: pageable host frame
    hb = B.Buffer.from_numpy(fr[0], W, H, B.kBandNV21)
    with pytest.raises(B.BandError):
        proc.record_batch([hb], [t])
    # refused: two frames whose tensors are separate allocations at no uniform stride -> pointer table
    t2 = B.Tensor(np.float32, (1, d, d, 3), device=True)
    t3 = B.Tensor(np.float32, (1, d, d, 3), device=True)
    dev3 = L.BandDeviceAlloc(3 * fb + 4096)
    bufs = [B.Buffer().set_raw(dev3 + o, W, H, B.kBandNV21, device=True) for o in (0, fb + 256, 2 * fb + 1024)]
    with pytest.raises(B.BandError):
        proc.record_batch(bufs, [t, t2, t3])
    assert proc.process(buf, t) == B.kBandOk      # the processor survives a refused recording
    L.BandDeviceFree(dev)
    L.BandDeviceFree(dev3)
