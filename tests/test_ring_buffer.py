"""band::TensorRingBuffer with host or HBM slots (reference band/tensor_ring_buffer.cc:27-161).
CPU tests: handle arithmetic and validity rules of the reference on a host ring. GPU tests: device slots are one
slab, the image processor writes model input straight into a lent slot, Put/Get cross memory spaces."""
import numpy as np
import pytest

from band_b200 import band_c as B


def _host_tensor(dtype, dims, fill=None):
    t = B.Tensor(dtype, dims, device=False)
    if fill is not None:
        import ctypes as C
        a = np.ascontiguousarray(fill, dtype=dtype)
        C.memmove(t.data_ptr, a.ctypes.data, a.nbytes)
    return t


def test_host_ring_handles_and_round_trip():
    a = _host_tensor(np.float32, (1, 4, 4, 3))
    b = _host_tensor(np.uint8, (1, 10))
    ring = B.TensorRingBuffer([a, b], [7, 9], size=4, device=False)
    assert not ring.is_handle_valid(0)          # nothing allocated yet (head_ == 0)
    hs = [ring.alloc() for _ in range(3)]
    assert hs == [0, 1, 2]
    assert all(ring.is_handle_valid(h) for h in hs) and not ring.is_handle_valid(3) and not ring.is_handle_valid(-1)
    rng = np.random.default_rng(5)
    vals = {}
    for h in hs:
        va = rng.standard_normal((1, 4, 4, 3)).astype(np.float32)
        vb = rng.integers(0, 256, size=(1, 10), dtype=np.uint8)
        sa, sb = _host_tensor(np.float32, (1, 4, 4, 3), va), _host_tensor(np.uint8, (1, 10), vb)
        assert ring.put(sa, 7, h) == B.kBandOk and ring.put(sb, 9, h) == B.kBandOk
        vals[h] = (va, vb)
    for h in hs:
        da, db = _host_tensor(np.float32, (1, 4, 4, 3)), _host_tensor(np.uint8, (1, 10))
        assert ring.get(da, 7, h) == B.kBandOk and ring.get(db, 9, h) == B.kBandOk
        assert np.array_equal(da.numpy(), vals[h][0]) and np.array_equal(db.numpy(), vals[h][1])
    # invalid tensor index / shape mismatch / stale handle -> kBandErr like the reference's InternalError
    assert ring.put(a, 8, 0) == B.kBandErr
    assert ring.put(b, 7, 0) == B.kBandErr      # u8 [1,10] into the f32 slot: CopyDataFrom refuses
    for _ in range(3):
        ring.alloc()                            # head_ = 6, size 4: handles 0 and 1 are overwritten
    assert not ring.is_handle_valid(1) and ring.is_handle_valid(2) and ring.is_handle_valid(5)
    assert ring.get(a, 7, 1) == B.kBandErr
    assert ring.slot(7, 1) is None and ring.slot(8, 5) is None
    # handle 5 shares ring index 1 with the stale handle 1 (handle % size)
    s5 = ring.slot(7, 5)
    assert s5 is not None and s5.nbytes == 4 * 4 * 3 * 4
    ring.close()


def test_ring_create_rejects_bad_arguments():
    import ctypes as C
    L = B.lib()
    assert not L.BandTensorRingBufferCreate(None, None, 0, 4, B.kBandHostMemory)
    a = _host_tensor(np.uint8, (2, 2))
    ta = (C.c_void_p * 1)(a.handle)
    ia = (C.c_int * 1)(0)
    assert not L.BandTensorRingBufferCreate(ta, ia, 1, 0, B.kBandHostMemory)
    assert L.BandTensorRingBufferAlloc(None) == -1
    assert L.BandTensorRingBufferGetSlot(None, 0, 0) is None


@pytest.mark.gpu
def test_device_ring_is_one_slab_and_processor_writes_into_slots(O):
    """Frames -> ColorSpaceConvert -> Resize -> Normalize land directly in HBM ring slots; the bytes a consumer
    reads back through GetTensorFromHandle equal the oracle's."""
    from gpu_util import to_capi_ops, rand_frames, oracle_chain, assert_same_bits
    NORM = (O.OP_NORMALIZE, 0, 0, 0, 0, 127.5, 127.5)
    W, H, d = 640, 480, 224
    fr = rand_frames(O, 6, W, H, O.FMT_NV21, 99)
    ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, d, d), NORM]
    st, want = oracle_chain(O, ops, fr, W, H, O.FMT_NV21, O.DT_F32, (1, d, d, 3))
    assert st == 0
    like = B.Tensor(np.float32, (1, d, d, 3), device=False)
    aux = B.Tensor(np.uint8, (1, 5), device=False)
    ring = B.TensorRingBuffer([like, aux], [0, 3], size=4, device=True)
    hs = [ring.alloc() for _ in range(4)]
    slots = [ring.slot(0, h) for h in hs]
    ptrs = [s.data_ptr for s in slots]
    assert all(p % 256 == 0 for p in ptrs)
    pitch = ptrs[1] - ptrs[0]
    assert pitch == (d * d * 3 * 4 + 255) // 256 * 256 + 256 and all(ptrs[i + 1] - ptrs[i] == pitch for i in range(3))
    proc = B.ImageProcessor(to_capi_ops(ops))
    bufs = [B.Buffer.from_numpy(fr[i], W, H, B.kBandNV21) for i in range(4)]
    assert proc.process_batch(bufs, slots) == B.kBandOk, B.last_error()
    for i, h in enumerate(hs):
        out = B.Tensor(np.float32, (1, d, d, 3), device=False)
        assert ring.get(out, 0, h) == B.kBandOk
        assert_same_bits(out.numpy(), want[i], "slot %d" % i)
        assert_same_bits(slots[i].numpy(), want[i], "slot view %d" % i)
    # wrap: handle 4 reuses ring index 0; single-frame Process into the lent slot; device -> device Get
    h4 = ring.alloc()
    assert not ring.is_handle_valid(hs[0])
    s4 = ring.slot(0, h4)
    assert s4.data_ptr == ptrs[0]
    assert proc.process(B.Buffer.from_numpy(fr[5], W, H, B.kBandNV21), s4) == B.kBandOk, B.last_error()
    dev_out = B.Tensor(np.float32, (1, d, d, 3), device=True)
    assert ring.get(dev_out, 0, h4) == B.kBandOk
    assert_same_bits(dev_out.numpy(), want[5], "wrapped slot")
    # Put from a host tensor into a device slot (H2D) and read it back
    src = _host_tensor(np.uint8, (1, 5), np.array([[1, 2, 3, 4, 5]], dtype=np.uint8))
    assert ring.put(src, 3, h4) == B.kBandOk
    back = B.Tensor(np.uint8, (1, 5), device=False)
    assert ring.get(back, 3, h4) == B.kBandOk and back.numpy().tolist() == [[1, 2, 3, 4, 5]]
    proc.close()
    ring.close()
