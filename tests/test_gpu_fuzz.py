"""Seeded slice of the differential fuzzer (scripts/fuzz_chains.py): random operator chains, formats, odd and tiny
dimensions, batches, host and HBM-resident buffers -- product (CUDA, through the C API) against the oracle, bit for bit,
and the same accept / reject verdict. The first run of the fuzzer found a kernel bug the hand-written cases had missed
(byte stores of the small-source up-scaler on rows that are not word-aligned); the seeds that hit it are pinned here."""
import os
import sys

import pytest

from conftest import ROOT

sys.path.insert(0, os.path.join(ROOT, "scripts"))


def _run(seeds):
    import fuzz_chains as F
    worker = F.OracleWorker()
    counts, fails = {}, []
    try:
        for s in seeds:
            c = F.make_case(s)
            v, detail = F.run_case(c, worker)
            counts[v] = counts.get(v, 0) + 1
            if v == "FAIL":
                fails.append("%s :: %s" % (F.describe(c), detail))
    finally:
        worker.close()
    return counts, fails


@pytest.mark.gpu
@pytest.mark.parametrize("first", [1, 401, 801, 1201])
def test_random_chains_match_the_oracle(O, first):
    counts, fails = _run(range(first, first + 400))
    assert not fails, "\n".join(fails[:10])
    assert counts.get("same", 0) >= 150, counts   # the generator must keep producing chains the reference accepts


def test_fuzz_generator_is_deterministic_and_mostly_valid(O):
    """CPU: the same seed gives the same case, and the oracle accepts a good share of the generated chains."""
    import fuzz_chains as F
    assert F.describe(F.make_case(77)) == F.describe(F.make_case(77))
    ok = rej = 0
    for s in range(1, 61):
        r = F.oracle_side(s)
        if r.get("shape") is not None or r.get("st", 1) == 0:
            ok += 1
        else:
            rej += 1
    assert ok >= 25, (ok, rej)


@pytest.mark.gpu
@pytest.mark.parametrize("first", [1, 61])
def test_random_sessions_match_a_persistent_oracle_processor(O, first):
    """scripts/fuzz_sessions.py: one processor, many calls -- refilled buffers, new handles, other layouts, other memory
    kinds, other batch sizes, host and HBM tensors -- against an oracle processor that lives as long."""
    import fuzz_sessions as S
    bad = []
    for s in range(first, first + 60):
        bad += S.run_session(s)
    assert not bad, "\n".join(bad[:10])


@pytest.mark.gpu
@pytest.mark.parametrize("first_op", ["flip", "rotate", "crop"])
def test_planar_layout_may_change_under_an_index_first_processor(O, first_op):
    """A processor whose first operator is Crop / Rotate / Flip caches an intermediate in the first frame's layout. The
    reference accepts the other planar layout on later calls and resolves U and V by name on both sides
    (buffer.cc:367-388, libyuv_image_operator.cc *Yv routines): same pixels as a fresh processor. Semi-planar pairs are
    refused (the reference exchanges U and V there, or writes the plane one byte off)."""
    import numpy as np
    from band_b200 import band_c as B
    from gpu_util import assert_same_bits, oracle_chain, rand_frames, to_capi_ops
    W, H = 96, 64
    pre = {"flip": (O.OP_FLIP, 1, 0), "rotate": (O.OP_ROTATE, 90), "crop": (O.OP_CROP, 8, 4, 71, 51)}[first_op]
    ow, oh = {"flip": (48, 32), "rotate": (32, 48), "crop": (32, 24)}[first_op]
    ops = [pre, (O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, ow, oh)]
    proc = B.ImageProcessor(to_capi_ops(ops))
    t = B.Tensor(np.uint8, (1, oh, ow, 3), device=False)
    for k, name in enumerate(["YV21", "YV12", "YV21", "YV12"]):
        fmt = getattr(O, "FMT_" + name)
        f = rand_frames(O, 1, W, H, fmt, seed=300 + k)
        st, want = oracle_chain(O, ops, f, W, H, fmt, O.DT_U8, (1, oh, ow, 3))
        assert st == 0
        buf = B.Buffer.from_numpy(f[0], W, H, B.FORMATS[name])
        assert proc.process(buf, t) == B.kBandOk, B.last_error()
        assert_same_bits(t.numpy(), want[0], "call %d (%s)" % (k, name))
        buf.close()
    # semi-planar after planar: refused, and the processor keeps working afterwards
    f = rand_frames(O, 1, W, H, O.FMT_NV12, seed=310)
    buf = B.Buffer.from_numpy(f[0], W, H, B.FORMATS["NV12"])
    assert proc.process(buf, t) != B.kBandOk
    assert "formats must match" in B.last_error()
    buf.close()
    f = rand_frames(O, 1, W, H, O.FMT_YV12, seed=311)
    st, want = oracle_chain(O, ops, f, W, H, O.FMT_YV12, O.DT_U8, (1, oh, ow, 3))
    buf = B.Buffer.from_numpy(f[0], W, H, B.FORMATS["YV12"])
    assert proc.process(buf, t) == B.kBandOk, B.last_error()
    assert_same_bits(t.numpy(), want[0], "after the refused call")
    buf.close(); t.close(); proc.close()


@pytest.mark.gpu
def test_seeds_that_found_bugs_stay_fixed(O):
    """16 / 46 / 284 / 482: packed_small_up_kernel byte stores on rows that are not word-aligned.
    19022: a Crop whose box starts at (0, 0) left identity index maps, and the Normalize behind it took the smaller view
    for materialised memory (linear index over the UNcropped rows)."""
    counts, fails = _run([16, 46, 284, 482, 19022])
    assert not fails, "\n".join(fails)


@pytest.mark.gpu
@pytest.mark.parametrize("box", [(0, 0, 1, 1), (0, 0, 4, 2), (0, 0, 7, 1), (0, 0, 2, 3)])
@pytest.mark.parametrize("out_dt_name,mean,std", [("DT_U8", 0.0, 1.0), ("DT_F32", 127.5, 127.5), ("DT_I8", 128.0, 1.0)])
def test_crop_at_the_origin_behind_a_resize_then_normalize(O, box, out_dt_name, mean, std):
    from gpu_util import assert_same_bits, gpu_chain, oracle_chain, rand_frames
    W, H = 20, 6
    out_dt = getattr(O, out_dt_name)
    ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, 8, 4), (O.OP_CROP,) + box, (O.OP_NORMALIZE, 0, 0, 0, 0, mean, std)]
    dims = (1, box[3] - box[1] + 1, box[2] - box[0] + 1, 3)
    fr = rand_frames(O, 2, W, H, O.FMT_NV21, seed=77)
    st, want = oracle_chain(O, ops, fr, W, H, O.FMT_NV21, out_dt, dims)
    assert st == 0
    for resident in (False, True):
        st, got = gpu_chain(ops, fr, W, H, O.FMT_NV21, out_dt, dims, device_resident=resident)
        assert st == 0
        assert_same_bits(got, want, "crop %s -> %s" % (box, out_dt_name))
