"""bandcu_two_stage_run_noalloc (include/bandcu.h): the two-stage chain whose intermediate image lives in a file-scope
device array instead of cudaMalloc'ed scratch. No counterpart in the reference -- added
for callers that cannot allocate device memory.

Checked through the C-ABI against the EXISTING entry points on the same inputs, bit for bit (u8 results: the tolerance
of every integer parity test here is zero), on two calls in a row -- the second with another frame size, so that the
array still holds the first call's bytes -- and against the full C-API chain (BandImageProcessor: ColorSpaceConvert ->
Resize), which the parity suite ties to the oracle.

Collected last on purpose (a failure here must not hide other tests)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SRC_RGBA, SRC_YUV420 = 2, 3
CVT_YUV2RGBA, CVT_RGBA2RGB = 2, 6
FLAVOR_ARGB, SCALE_COPY, SCALE_BILINEAR_UP, DT_U8 = 0, 0, 3, 3
E_INVALID = 10001


def _fixed_div1(num, div):  # scale_common.cc FixedDiv1
    return ((num << 16) - 0x00010001) // (div - 1)


def _descs(S, W, H, DW, DH):
    """NV12 W x H -> RGBA (same size), then RGBA -> RGB bilinear up-scale to DW x DH: what the plan compiler emits for
    ColorSpaceConvert(RGB) -> Resize(up) on a 4:2:0 source. Slopes: scale_common.cc ScaleSlope, kFilterBilinear,
    dst > src on both axes (band_b200/csrc/band/buffer/scale_dispatch.cc DispatchPacked)."""
    ident = S.Affine(0, 1, 0, 0, 0, 1)
    a = S.PlanDesc()
    a.src_kind, a.stride0, a.stride_uv, a.pixel_stride_uv = SRC_YUV420, W, W, 2
    a.vw, a.vh = W, H
    a.luma_map = a.pre_map = a.chroma_map = ident
    a.cvt, a.flavor, a.scale_mode = CVT_YUV2RGBA, FLAVOR_ARGB, SCALE_COPY
    a.dx = a.dy = 0x10000
    a.dst_width, a.dst_height, a.channels, a.out_dtype = W, H, 4, DT_U8
    a.std_ = 1.0
    a.dst_row_stride = W * 4
    a.src_width, a.src_height = W, H
    b = S.PlanDesc()
    b.src_kind, b.stride0 = SRC_RGBA, W * 4
    b.vw, b.vh = W, H
    b.luma_map = b.pre_map = b.chroma_map = ident
    b.cvt, b.flavor, b.scale_mode, b.linear = CVT_RGBA2RGB, FLAVOR_ARGB, SCALE_BILINEAR_UP, 0
    b.x0 = b.y0 = 0
    b.dx, b.dy = _fixed_div1(W, DW), _fixed_div1(H, DH)
    b.max_y = (H - 1) << 16
    b.dst_width, b.dst_height, b.channels, b.out_dtype = DW, DH, 3, DT_U8
    b.std_ = 1.0
    b.dst_row_stride = DW * 3
    b.src_width, b.src_height = W, H
    return a, b


class _Dev:
    def __init__(self, S, nbytes):
        self.S, self.p = S, C.c_void_p()
        S.check(S.lib().bandcu_malloc(C.byref(self.p), nbytes), "bandcu_malloc")

    def free(self):
        self.S.lib().bandcu_free(self.p)


def _existing(S, a, b, frame_dev, W, H, out_dev):
    """The existing entry points: one plan per stage, cudaMalloc'ed scratch between them."""
    L = S.lib()
    pa, pb = C.c_void_p(), C.c_void_p()
    S.check(L.bandcu_plan_create(C.byref(a), C.byref(pa)), "plan_create first")
    S.check(L.bandcu_plan_create(C.byref(b), C.byref(pb)), "plan_create second")
    scratch = _Dev(S, W * H * 4)
    try:
        f1 = S.Frame(frame_dev, frame_dev + W * H, frame_dev + W * H + 1, scratch.p.value)
        S.check(L.bandcu_plan_run_strided(pa, C.byref(f1), 0, 0, 1, None), "plan_run first")
        f2 = S.Frame(scratch.p.value, None, None, out_dev)
        S.check(L.bandcu_plan_run_strided(pb, C.byref(f2), 0, 0, 1, None), "plan_run second")
        S.check(L.bandcu_stream_sync(None), "sync")
    finally:
        scratch.free()
        L.bandcu_plan_destroy(pa)
        L.bandcu_plan_destroy(pb)


def _download(S, dev_ptr, nbytes):
    out = np.empty(nbytes, dtype=np.uint8)
    S.check(S.lib().bandcu_memcpy_d2h(out.ctypes.data, dev_ptr, nbytes, None), "d2h")
    S.check(S.lib().bandcu_stream_sync(None), "sync")
    return out


def test_no_alloc_entry_point_matches_the_plan_api_twice_in_a_row():
    from band_b200 import band_c as B
    from band_b200 import cuda_shim as S
    L = S.lib()
    assert S.NOALLOC_SCRATCH_BYTES == 8 << 20
    # second call: a smaller frame, so the array still holds most of the first call's image
    for call, (W, H, DW, DH) in enumerate([(640, 360, 1000, 600), (322, 182, 700, 401)]):
        rng = np.random.default_rng(900 + call)
        frame = rng.integers(0, 256, size=W * H + 2 * ((W + 1) // 2) * ((H + 1) // 2), dtype=np.uint8)
        a, b = _descs(S, W, H, DW, DH)
        src, want_dev, got_dev = _Dev(S, frame.size), _Dev(S, DW * DH * 3), _Dev(S, DW * DH * 3)
        try:
            S.check(L.bandcu_memcpy_h2d(src.p, frame.ctypes.data, frame.size, None), "h2d")
            S.check(L.bandcu_memset(want_dev.p, 0x5A, DW * DH * 3, None), "memset")
            S.check(L.bandcu_memset(got_dev.p, 0xA5, DW * DH * 3, None), "memset")
            _existing(S, a, b, src.p.value, W, H, want_dev.p.value)
            n0 = L.bandcu_launch_count()
            f = S.Frame(src.p.value, src.p.value + W * H, src.p.value + W * H + 1, got_dev.p.value)
            S.check(L.bandcu_two_stage_run_noalloc(C.byref(a), C.byref(b), C.byref(f), None), "two_stage_run_noalloc")
            assert L.bandcu_launch_count() - n0 == 2   # two of OUR kernels, nothing else
            want, got = _download(S, want_dev.p, DW * DH * 3), _download(S, got_dev.p, DW * DH * 3)
            bad = int((want != got).sum())
            assert bad == 0, "call %d: %d of %d bytes differ from the plan API" % (call, bad, want.size)
            # and the public chain a user of Band's API would run (tied to the oracle by the parity suite)
            ops = [(B.BAND_COLOR_SPACE_CONVERT, B.FORMATS["RGB"]), (B.BAND_RESIZE, DW, DH)]
            st, chain = B.run_chain(ops, frame[None], W, H, B.FORMATS["NV12"], np.uint8, (1, DH, DW, 3), device_resident=True)
            assert st == B.kBandOk, B.last_error()
            bad = int((chain.reshape(-1) != got).sum())
            assert bad == 0, "call %d: %d of %d bytes differ from BandImageProcessor's chain" % (call, bad, got.size)
        finally:
            for d in (src, want_dev, got_dev):
                d.free()


def test_no_alloc_entry_point_rejects_what_does_not_fit():
    from band_b200 import cuda_shim as S
    L = S.lib()
    W, H = 2048, 1100   # 9 011 200 bytes of RGBA: above the 8 MiB array
    a, b = _descs(S, W, H, 2100, 1200)
    dummy = _Dev(S, 4096)
    try:
        f = S.Frame(dummy.p.value, dummy.p.value, dummy.p.value, dummy.p.value)
        n0 = L.bandcu_launch_count()
        assert L.bandcu_two_stage_run_noalloc(C.byref(a), C.byref(b), C.byref(f), None) == E_INVALID
        msg = L.bandcu_last_error_string().decode()
        assert "BANDCU_NOALLOC_SCRATCH_BYTES" in msg and "9011200" in msg, msg
        # a second stage that reads outside the first stage's image is refused too
        a, b = _descs(S, 64, 48, 128, 96)
        b.vw = 65
        assert L.bandcu_two_stage_run_noalloc(C.byref(a), C.byref(b), C.byref(f), None) == E_INVALID
        assert "index map" in L.bandcu_last_error_string().decode()
        # a first stage that does not leave u8 pixels
        a, b = _descs(S, 64, 48, 128, 96)
        a.normalize = 1
        assert L.bandcu_two_stage_run_noalloc(C.byref(a), C.byref(b), C.byref(f), None) == E_INVALID
        assert L.bandcu_two_stage_run_noalloc(None, C.byref(b), C.byref(f), None) == E_INVALID
        assert L.bandcu_launch_count() == n0   # nothing was launched
    finally:
        dummy.free()
