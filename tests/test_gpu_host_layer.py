"""Host-layer behaviour around the fused plan (no reference counterpart: the reference has no device, no batches and
no graphs). What must hold is stated against the reference all the same:

  * a chain writes `width` bytes per output row and nothing else, like libyuv's row functions behind
    BufferProcessor::Process (/root/reference/band/buffer/buffer_processor.cc:22-43): padding of a strided host
    output, the gaps between its planes and the bytes between pixel-strided samples stay the caller's;
  * a recorded batch keeps working whatever the processor does afterwards;
  * one call shards a batch over devices and gives the bits a single processor gives;
  * an unmodified caller's pageable frames (BandBufferSetFromRawData, c_api_buffer.cc:40-53) reach the in-place PCIe
    read path through the registration cache, bit-identical.
"""
import ctypes as C

import numpy as np
import pytest

from gpu_util import assert_same_bits, oracle_chain, rand_frames, to_capi_ops

NORM = (5, 0, 0, 0, 0, 127.5, 127.5)
SENT = 0xA5


# ---------------------------------------------------------------------------------------------------------------
# CPU: the slicing rule (no device needed)
# ---------------------------------------------------------------------------------------------------------------
def test_shard_slices_are_contiguous_and_cover_the_batch():
    from band_b200 import band_c as B
    from band_b200 import shard
    for n in (0, 1, 7, 24, 512, 4096, 4097):
        for world in (1, 2, 3, 4, 8):
            covered = []
            for r in range(world):
                b, e = B.shard_slice(n, world, r)
                assert (b, e) == shard.shard_slice(n, world, r)
                assert 0 <= b <= e <= n
                covered.extend(range(b, e))
            assert covered == list(range(n)), (n, world)
    assert B.shard_slice(10, 0, 0) == (0, 0)
    assert B.shard_slice(10, 2, 5) == (0, 0)


# ---------------------------------------------------------------------------------------------------------------
# GPU
# ---------------------------------------------------------------------------------------------------------------
def _planes_of(frame, W, H, fmt_name):
    cw, ch = (W + 1) // 2, (H + 1) // 2
    Y = frame[:W * H].reshape(H, W)
    c = frame[W * H:]
    if fmt_name in ("NV12", "NV21"):
        pair = c.reshape(ch, cw, 2)
        U, V = (pair[..., 0], pair[..., 1]) if fmt_name == "NV12" else (pair[..., 1], pair[..., 0])
    else:
        a, b = c[:cw * ch].reshape(ch, cw), c[cw * ch:].reshape(ch, cw)
        U, V = (a, b) if fmt_name == "YV21" else (b, a)
    return Y, U, V


@pytest.mark.gpu
@pytest.mark.parametrize("fmt_name,layout", [("NV12", "interleaved"), ("NV21", "interleaved"), ("YV12", "planar"),
                                             ("NV12", "split_ps2")])
@pytest.mark.parametrize("op", [(3, 1, 0), (2, 180), (0, 32, 16, 351, 255)], ids=["flip", "rot180", "crop"])
def test_padded_host_output_keeps_the_callers_padding(O, fmt_name, layout, op):
    """A 4:2:0 -> 4:2:0 chain into a HOST image whose rows are padded, whose planes have gaps between them and (split
    layout) whose chroma samples sit every second byte of two separate buffers: every byte the chain does not own must
    still hold the caller's sentinel afterwards, and the owned bytes must equal the oracle."""
    from band_b200 import band_c as B
    W, H = 640, 480
    fmt = getattr(O, "FMT_" + fmt_name)
    frame = rand_frames(O, 1, W, H, fmt, seed=11)[0]
    st, want_raw, (ow, oh, ofmt) = O.run_op(op, frame, W, H, fmt)
    assert st == 0 and ofmt == fmt
    wY, wU, wV = _planes_of(want_raw, ow, oh, fmt_name)
    cw, ch = (ow + 1) // 2, (oh + 1) // 2
    ys = ow + 40
    # one backing block, planes separated by gaps
    gap = 1000
    if layout == "planar":
        cs, ps = cw + 24, 1
    else:
        cs, ps = 2 * cw + 24, 2
    y_off = 64
    c1_off = y_off + ys * oh + gap
    c2_off = c1_off + cs * ch + gap
    total = c2_off + cs * ch + gap
    back = np.full(total, SENT, dtype=np.uint8)
    if layout == "interleaved":
        first_is_u = fmt_name == "NV12"
        u_off, v_off = (c1_off, c1_off + 1) if first_is_u else (c1_off + 1, c1_off)
    else:
        u_off, v_off = c1_off, c2_off
    base = back.ctypes.data
    out_buf = B.Buffer().set_yuv(base + y_off, base + u_off, base + v_off, ow, oh, ys, cs, ps, B.FORMATS[fmt_name],
                                 keep=back)
    in_buf = B.Buffer.from_numpy(frame, W, H, B.FORMATS[fmt_name])
    proc = B.ImageProcessor(to_capi_ops([op]))
    ia, oa = (C.c_void_p * 1)(in_buf.handle), (C.c_void_p * 1)(out_buf.handle)
    assert B.lib().BandImageProcessorProcessBuffers(proc.handle, ia, oa, 1, 0) == B.kBandOk, B.last_error()

    owned = np.zeros(total, dtype=bool)
    yv = back[y_off:y_off + ys * oh].reshape(oh, ys)
    assert np.array_equal(yv[:, :ow], wY), "luma"
    owned[y_off:y_off + ys * oh].reshape(oh, ys)[:, :ow] = True
    for off, want in ((u_off, wU), (v_off, wV)):
        rows = np.lib.stride_tricks.as_strided(back[off:], shape=(ch, cw), strides=(cs, ps))
        assert np.array_equal(rows, want), "chroma"
        np.lib.stride_tricks.as_strided(owned[off:], shape=(ch, cw), strides=(cs, ps))[:] = True
    untouched = back[~owned]
    assert (untouched == SENT).all(), "%d bytes outside the image were overwritten" % int((untouched != SENT).sum())
    for x in (in_buf, out_buf, proc):
        x.close()


@pytest.mark.gpu
def test_recorded_batch_survives_plan_changes_and_processor_deletion(O):
    """the graph reads plan-owned device memory (Normalize table). Feeding the processor another layout
    (which compiles and caches another plan) or deleting it must not pull that memory from under a replay."""
    from band_b200 import band_c as B
    L = B.lib()
    W, H, dw, dh = 640, 480, 224, 224
    ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, dw, dh), NORM]
    frames = rand_frames(O, 2, W, H, O.FMT_NV21, seed=5)
    st, want = oracle_chain(O, ops, frames, W, H, O.FMT_NV21, O.DT_F32, (1, dh, dw, 3))
    assert st == 0
    fb = frames.shape[1]
    pitch = (fb + 255) // 256 * 256
    dev = L.BandDeviceAlloc(pitch * 2)
    for i in range(2):
        assert L.BandCopyToDevice(dev + i * pitch, frames[i].ctypes.data, fb) == B.kBandOk
    bufs = [B.Buffer().set_raw(dev + i * pitch, W, H, B.kBandNV21, device=True) for i in range(2)]
    # recorded batches of more than one frame need uniformly strided tensors: carve them out of ring slots is overkill
    # here, so record one frame at a time
    tens = [B.Tensor(np.float32, (1, dh, dw, 3), device=True) for _ in range(2)]
    proc = B.ImageProcessor(to_capi_ops(ops))
    recs = [proc.record_batch([bufs[i]], [tens[i]]) for i in range(2)]
    # other layouts through the same processor, more than the plan cache holds. Same dimensions throughout: like the
    # reference, a processor whose first operator caches its intermediate is bound to the first frame size
    # (ColorSpaceConvert validates formats only and the conversion then rejects the stale dimensions,
    # libyuv_image_operator.cc:1709-1727) -- what changes between cameras here is the 4:2:0 layout and the strides.
    others = [("NV12", 0), ("YV12", 0), ("YV21", 0), ("NV12", 64), ("YV12", 32)]
    for k, (fmt_name, pad) in enumerate(others):
        fmt = getattr(O, "FMT_" + fmt_name)
        other = rand_frames(O, 1, W, H, fmt, seed=20 + k)
        st2, want2 = oracle_chain(O, ops, other, W, H, fmt, O.DT_F32, (1, dh, dw, 3))
        assert st2 == 0
        t2 = B.Tensor(np.float32, (1, dh, dw, 3), device=False)
        if pad == 0:
            b2 = B.Buffer.from_numpy(other[0], W, H, B.FORMATS[fmt_name])
        else:  # the same pixels with padded rows (BandBufferSetFromYUVData): another plan key
            Y, U, V = _planes_of(other[0], W, H, fmt_name)
            cw = (W + 1) // 2
            yb = np.zeros((H, W + pad), np.uint8); yb[:, :W] = Y
            if fmt_name in ("NV12", "NV21"):
                cb = np.zeros((U.shape[0], 2 * cw + pad), np.uint8)
                cb[:, 0:2 * cw:2], cb[:, 1:2 * cw:2] = (U, V) if fmt_name == "NV12" else (V, U)
                up, vp = (cb.ctypes.data, cb.ctypes.data + 1) if fmt_name == "NV12" else (cb.ctypes.data + 1, cb.ctypes.data)
                keep, cs, ps = (yb, cb), 2 * cw + pad, 2
            else:
                ub = np.zeros((U.shape[0], cw + pad), np.uint8); ub[:, :cw] = U
                vb = np.zeros((U.shape[0], cw + pad), np.uint8); vb[:, :cw] = V
                up, vp, keep, cs, ps = ub.ctypes.data, vb.ctypes.data, (yb, ub, vb), cw + pad, 1
            b2 = B.Buffer().set_yuv(yb.ctypes.data, up, vp, W, H, W + pad, cs, ps, B.FORMATS[fmt_name], keep=keep)
        assert proc.process(b2, t2) == B.kBandOk, B.last_error()
        assert_same_bits(t2.numpy(), want2[0], "other layout %s pad %d" % (fmt_name, pad))
        b2.close(); t2.close()
    for i in range(2):
        assert L.BandCopyToDevice(tens[i].data_ptr, np.zeros(tens[i].nbytes, np.uint8).ctypes.data, tens[i].nbytes) == B.kBandOk
        assert B.ImageProcessor.replay(recs[i]) == B.kBandOk, B.last_error()
        assert_same_bits(tens[i].numpy(), want[i], "replay after plan changes")
    proc.close()  # the recordings keep their plan
    for i in range(2):
        assert L.BandCopyToDevice(tens[i].data_ptr, np.zeros(tens[i].nbytes, np.uint8).ctypes.data, tens[i].nbytes) == B.kBandOk
        assert B.ImageProcessor.replay(recs[i]) == B.kBandOk, B.last_error()
        assert_same_bits(tens[i].numpy(), want[i], "replay after the processor is gone")
        B.ImageProcessor.release(recs[i])
    for x in bufs + tens:
        x.close()
    L.BandDeviceFree(dev)


@pytest.mark.gpu
def test_recording_a_staged_batch_is_refused_cleanly(O):
    """Pinned host frames are accepted by RecordBatch only when the whole pass is one in-place launch. A chain that
    runs on the tile-staged kernel stages host frames through copies: the call must fail with a message (not with an
    invalidated capture) and the processor must keep working."""
    from band_b200 import band_c as B
    L = B.lib()
    W, H = 640, 480
    ops = [(O.OP_ROTATE, 90), (O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, 120, 160), NORM]
    frame = rand_frames(O, 1, W, H, O.FMT_NV21, seed=9)
    st, want = oracle_chain(O, ops, frame, W, H, O.FMT_NV21, O.DT_F32, (1, 160, 120, 3))
    assert st == 0
    host = L.BandHostAlloc(frame.nbytes)
    C.memmove(host, frame.ctypes.data, frame.nbytes)
    buf = B.Buffer().set_raw(host, W, H, B.kBandNV21)
    t = B.Tensor(np.float32, (1, 160, 120, 3), device=False)
    proc = B.ImageProcessor(to_capi_ops(ops))
    with pytest.raises(B.BandError) as err:
        proc.record_batch([buf], [t])
    assert "staging" in str(err.value), str(err.value)
    assert proc.process(buf, t) == B.kBandOk, B.last_error()
    assert_same_bits(t.numpy(), want[0], "after a refused recording")
    for x in (buf, t, proc):
        x.close()
    L.BandHostFree(host)


@pytest.mark.gpu
def test_alternating_layouts_reuse_their_plans(O):
    """a one-entry plan cache re-lowered (cudaMalloc + table kernel + device sync) on every call of a
    processor fed two layouts in turn. Now each layout costs its set-up once: later calls are one launch each."""
    from band_b200 import band_c as B
    from band_b200 import hotpath
    ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, 224, 224), NORM]
    proc = B.ImageProcessor(to_capi_ops(ops))
    cases = []
    w, h = 640, 480  # one frame size (see the note on reference semantics above), two 4:2:0 layouts
    for k, fmt_name in enumerate(["NV12", "YV12"]):
        fmt = getattr(O, "FMT_" + fmt_name)
        f = rand_frames(O, 1, w, h, fmt, seed=40 + k)
        st, want = oracle_chain(O, ops, f, w, h, fmt, O.DT_F32, (1, 224, 224, 3))
        assert st == 0
        cases.append((B.Buffer.from_numpy(f[0], w, h, B.FORMATS[fmt_name]), want[0]))
    t = B.Tensor(np.float32, (1, 224, 224, 3), device=False)
    per_call = []
    for rnd in range(3):
        for buf, want in cases:
            n0 = hotpath.launch_count()
            assert proc.process(buf, t) == B.kBandOk, B.last_error()
            per_call.append(hotpath.launch_count() - n0)
            assert_same_bits(t.numpy(), want, "round %d" % rnd)
    assert min(per_call[:2]) >= 2, per_call      # Normalize-table kernels + the fused kernel, once per layout
    assert per_call[2:] == [1, 1, 1, 1], per_call
    for buf, _ in cases:
        buf.close()
    t.close(); proc.close()


@pytest.mark.gpu
def test_processors_with_different_rings_share_a_kernel(O):
    """Two processors whose chains pick the SAME row-staged kernel with DIFFERENT shared-memory rings (1920- and
    640-wide NV12 sources), called in turn. The dynamic shared-memory limit is an attribute of the kernel, not of a plan:
    the wide plan's cached launch geometry must survive the narrow plan's set-up (the round-2 final bench crashed on
    exactly this with cudaErrorLaunchOutOfResources)."""
    from band_b200 import band_c as B
    ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, 224, 224), NORM]
    made = []
    for k, (w, h) in enumerate([(1920, 1080), (640, 480)]):
        f = rand_frames(O, 1, w, h, O.FMT_NV12, seed=70 + k)
        st, want = oracle_chain(O, ops, f, w, h, O.FMT_NV12, O.DT_F32, (1, 224, 224, 3))
        assert st == 0
        made.append((B.ImageProcessor(to_capi_ops(ops)), B.Buffer.from_numpy(f[0], w, h, B.FORMATS["NV12"]), want[0]))
    t = B.Tensor(np.float32, (1, 224, 224, 3), device=False)
    for rnd in range(3):
        for proc, buf, want in made:
            assert proc.process(buf, t) == B.kBandOk, B.last_error()
            assert_same_bits(t.numpy(), want, "round %d" % rnd)
    # and the other order of first use
    for proc, buf, _ in made:
        buf.close(); proc.close()
    made2 = []
    for k, (w, h) in enumerate([(640, 480), (1920, 1080)]):
        f = rand_frames(O, 1, w, h, O.FMT_NV12, seed=80 + k)
        st, want = oracle_chain(O, ops, f, w, h, O.FMT_NV12, O.DT_F32, (1, 224, 224, 3))
        made2.append((B.ImageProcessor(to_capi_ops(ops)), B.Buffer.from_numpy(f[0], w, h, B.FORMATS["NV12"]), want[0]))
    for rnd in range(2):
        for proc, buf, want in made2:
            assert proc.process(buf, t) == B.kBandOk, B.last_error()
            assert_same_bits(t.numpy(), want, "second order, round %d" % rnd)
    for proc, buf, _ in made2:
        buf.close(); proc.close()
    t.close()


@pytest.mark.gpu
@pytest.mark.parametrize("mem", ["host", "pinned", "device"])
def test_sharded_batch_equals_single_processor_and_oracle(O, mem):
    """24 frames through BandImageProcessorProcessBatchSharded == the same frames through one processor == the oracle.
    On a one-GPU box the shards are [0, 0, 0] (three processors, three host threads, one device); with more GPUs
    every GPU is a shard and device-resident slices are placed on their shard's device."""
    from band_b200 import band_c as B
    L = B.lib()
    n, W, H, dw, dh = 24, 640, 480, 224, 224
    ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, dw, dh), NORM]
    frames = rand_frames(O, n, W, H, O.FMT_NV12, seed=77)
    st, want = oracle_chain(O, ops, frames, W, H, O.FMT_NV12, O.DT_F32, (1, dh, dw, 3))
    assert st == 0
    ndev = L.BandCudaGetDeviceCount()
    devices = list(range(ndev)) if ndev > 1 else [0, 0, 0]
    sp = B.ShardedImageProcessor(to_capi_ops(ops), devices)
    assert sp.num_shards == len(devices)
    fb = frames.shape[1]
    bufs, tens, frees = [None] * n, [None] * n, []
    for k in range(sp.num_shards):
        b, e = sp.slice_of(n, k)
        assert (b, e) == B.shard_slice(n, len(devices), k)
        assert L.BandCudaSetDevice(sp.device_of(k)) == B.kBandOk
        for i in range(b, e):
            if mem == "device":
                d = L.BandDeviceAlloc(fb)
                assert d and L.BandCopyToDevice(d, frames[i].ctypes.data, fb) == B.kBandOk
                frees.append(("dev", d))
                bufs[i] = B.Buffer().set_raw(d, W, H, B.kBandNV12, device=True)
            elif mem == "pinned":
                h = L.BandHostAlloc(fb)
                C.memmove(h, frames[i].ctypes.data, fb)
                frees.append(("host", h))
                bufs[i] = B.Buffer().set_raw(h, W, H, B.kBandNV12)
            else:
                bufs[i] = B.Buffer.from_numpy(frames[i], W, H, B.kBandNV12)
            tens[i] = B.Tensor(np.float32, (1, dh, dw, 3), device=(mem == "device"))
    L.BandCudaSetDevice(0)
    assert sp.process_batch(bufs, tens) == B.kBandOk, B.last_error()
    got = np.stack([t.numpy() for t in tens])
    assert_same_bits(got, want, "sharded (%s) vs oracle" % mem)
    # asynchronous form, two batches queued back to back, then one synchronize
    for t in tens:
        if mem != "device":
            C.memset(t.data_ptr, 0, t.nbytes)
    ba = (C.c_void_p * n)(*[b.handle for b in bufs])
    ta = (C.c_void_p * n)(*[t.handle for t in tens])
    half = n // 2
    if mem != "device":
        ba2 = (C.c_void_p * (n - half))(*[b.handle for b in bufs[half:]])
        ta2 = (C.c_void_p * (n - half))(*[t.handle for t in tens[half:]])
        assert sp.process_batch_async(ba, ta, half) == B.kBandOk
        assert sp.process_batch_async(ba2, ta2, n - half) == B.kBandOk
    else:  # device slices belong to one shard each: a different cut would cross devices
        for t in tens:
            assert L.BandCopyToDevice(t.data_ptr, np.zeros(t.nbytes, np.uint8).ctypes.data, t.nbytes) == B.kBandOk
        assert sp.process_batch_async(ba, ta, n) == B.kBandOk
    assert sp.synchronize() == B.kBandOk, B.last_error()
    assert_same_bits(np.stack([t.numpy() for t in tens]), want, "sharded async (%s)" % mem)
    if mem != "device":
        st1, single = B.run_chain(to_capi_ops(ops), frames, W, H, B.kBandNV12, np.float32, (1, dh, dw, 3))
        assert st1 == B.kBandOk
        assert_same_bits(got, single, "sharded vs one processor")
    # errors surface from the shard that hit them
    bad = B.Tensor(np.float32, (1, dh + 1, dw, 3), device=(mem == "device"))
    assert sp.process_batch(bufs[:1], [bad]) == B.kBandErr
    if mem == "device" and len(set(devices)) > 1:
        # a device-resident frame handed to the wrong shard is refused before anything launches (frame 1 of a 2-frame
        # batch belongs to shard 1, but bufs[1] lives on shard 0's GPU)
        assert sp.process_batch(bufs[:2], tens[:2]) == B.kBandErr
        assert "lives on device" in B.last_error(), B.last_error()
        assert sp.process_batch(bufs, tens) == B.kBandOk, "a refused batch must not poison the next one"
    else:
        assert sp.process_batch(bufs[:2], tens[:2]) == B.kBandOk, "a failed batch must not poison the next one"
    bad.close()
    sp.close()
    for x in bufs + tens:
        x.close()
    for kind, p in frees:
        (L.BandDeviceFree if kind == "dev" else L.BandHostFree)(p)


@pytest.mark.gpu
def test_registration_cache_brings_pageable_frames_to_the_in_place_path(O):
    """64 pageable (numpy / malloc) frames: without the cache they are staged in 32-frame chunks (two fused launches);
    with it they are registered on first sight and read in place (one launch), bit-identical, found again on the
    second call (no new registrations), and unregistered when the processor goes away."""
    from band_b200 import band_c as B
    from band_b200 import hotpath
    n, W, H, dw, dh = 64, 640, 480, 224, 224
    ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, dw, dh), NORM]
    # a "camera ring": one big malloc'ed block, frames at a 16-byte granular pitch, deliberately not page aligned
    fb = O.frame_bytes(W, H, O.FMT_NV12)
    block = np.zeros(n * fb + 4096 + 16, dtype=np.uint8)
    off = (-block.ctypes.data) % 4096 + 16
    ring = block[off:off + n * fb].reshape(n, fb)
    ring[:] = rand_frames(O, n, W, H, O.FMT_NV12, seed=123)
    st, want = oracle_chain(O, ops, ring, W, H, O.FMT_NV12, O.DT_F32, (1, dh, dw, 3))
    assert st == 0

    def run(proc):
        bufs = [B.Buffer().set_raw(ring[i].ctypes.data, W, H, B.kBandNV12, keep=block) for i in range(n)]
        tens = [B.Tensor(np.float32, (1, dh, dw, 3), device=False) for _ in range(n)]
        n0 = hotpath.launch_count()
        assert proc.process_batch(bufs, tens) == B.kBandOk, B.last_error()
        launches = hotpath.launch_count() - n0
        got = np.stack([t.numpy() for t in tens])
        for x in bufs + tens:
            x.close()
        return got, launches

    plain = B.ImageProcessor(to_capi_ops(ops))
    got0, l0 = run(plain)
    got0, l0 = run(plain)       # second call: the plan's table kernel is no longer in the count
    plain.close()
    assert_same_bits(got0, want, "staged pageable frames")
    assert l0 == 2, l0
    assert B.host_registration_cache(True, 1 << 30) == B.kBandOk
    try:
        cached = B.ImageProcessor(to_capi_ops(ops))
        got1, l1 = run(cached)
        s1 = B.host_registration_stats()
        got2, l2 = run(cached)
        s2 = B.host_registration_stats()
        assert_same_bits(got1, want, "registered frames, first sight")
        assert_same_bits(got2, want, "registered frames, second call")
        assert s1["registrations"] >= 1 and s1["registered_bytes"] >= n * fb, s1
        assert s2["registrations"] == s1["registrations"] and s2["hits"] > s1["hits"], (s1, s2)
        assert l2 == 1, "registered frames must run as one in-place launch, saw %d" % l2
        cached.close()
        s3 = B.host_registration_stats()
        assert s3["entries"] == 0 and s3["registered_bytes"] == 0, s3
    finally:
        B.host_registration_cache(False)


@pytest.mark.gpu
@pytest.mark.parametrize("mem", ["device", "pinned"])
def test_repeat_calls_replay_the_launch_and_follow_memory_and_handles(O, mem):
    """A camera loop hands the processor the SAME BandBuffer / BandTensor objects call after call: from the second call on
    the processor replays its launch without re-validating (Buffer::Epoch). The replay must read what the frame memory holds
    NOW, must notice a handle that was re-pointed (BandBufferSetFromRawData creates a new view), and a bad call in between
    must still fail."""
    from band_b200 import band_c as B
    from band_b200 import hotpath
    L = B.lib()
    n, W, H, dw, dh = 6, 640, 480, 224, 224
    ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, dw, dh), NORM]
    fa = rand_frames(O, n, W, H, O.FMT_NV12, seed=301)
    fb = rand_frames(O, n, W, H, O.FMT_NV12, seed=302)
    _, want_a = oracle_chain(O, ops, fa, W, H, O.FMT_NV12, O.DT_F32, (1, dh, dw, 3))
    _, want_b = oracle_chain(O, ops, fb, W, H, O.FMT_NV12, O.DT_F32, (1, dh, dw, 3))
    fbytes = fa.shape[1]
    pitch = (fbytes + 255) // 256 * 256
    dev = mem == "device"
    alloc, free = (L.BandDeviceAlloc, L.BandDeviceFree) if dev else (L.BandHostAlloc, L.BandHostFree)
    base, other = alloc(pitch * n), alloc(pitch * n)

    def fill(dst, frames):
        for i in range(n):
            if dev:
                assert L.BandCopyToDevice(dst + i * pitch, frames[i].ctypes.data, fbytes) == B.kBandOk
            else:
                C.memmove(dst + i * pitch, frames[i].ctypes.data, fbytes)

    fill(base, fa)
    fill(other, fb)
    bufs = [B.Buffer().set_raw(base + i * pitch, W, H, B.kBandNV12, device=dev) for i in range(n)]
    tens = [B.Tensor(np.float32, (1, dh, dw, 3), device=dev) for _ in range(n)]
    proc = B.ImageProcessor(to_capi_ops(ops))
    got = lambda: np.stack([t.numpy() for t in tens])
    assert proc.process_batch(bufs, tens) == B.kBandOk, B.last_error()
    assert_same_bits(got(), want_a, "first call")
    for _ in range(3):   # replays
        n0 = hotpath.launch_count()
        assert proc.process_batch(bufs, tens) == B.kBandOk, B.last_error()
        assert hotpath.launch_count() - n0 == 1
    assert_same_bits(got(), want_a, "replayed call")
    fill(base, fb)       # the camera refills the same memory
    assert proc.process_batch(bufs, tens) == B.kBandOk
    assert_same_bits(got(), want_b, "replay reads the refilled frames")
    fill(base, fa)
    bufs[2].set_raw(other + 2 * pitch, W, H, B.kBandNV12, device=dev)   # handle 2 now views another frame (holds fb[2])
    assert proc.process_batch(bufs, tens) == B.kBandOk
    want_mix = want_a.copy(); want_mix[2] = want_b[2]
    assert_same_bits(got(), want_mix, "a re-pointed handle is noticed")
    bad = B.Tensor(np.float32, (1, dh + 1, dw, 3), device=dev)
    assert proc.process_batch(bufs[:1], [bad]) == B.kBandErr
    assert proc.process_batch(bufs, tens) == B.kBandOk
    assert_same_bits(got(), want_mix, "after a failed call")
    # single-frame entry point, repeated
    for _ in range(3):
        assert proc.process(bufs[0], tens[0]) == B.kBandOk
    assert_same_bits(tens[0].numpy(), want_a[0], "BandImageProcessorProcess repeated")
    for x in bufs + tens + [bad, proc]:
        x.close()
    free(base); free(other)
