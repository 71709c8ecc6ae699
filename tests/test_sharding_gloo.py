"""N>1 host logic on CPU: world_size-2 gloo process group. Each rank takes its shard of a seeded batch, runs it
(through the oracle here -- there is no GPU in this container; the GPU box runs the same sharding with the CUDA
path in bench.py), and the union of the shards must equal the single-process result; the timing reduction is
the max over ranks."""
import hashlib
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT

WORKER = r'''
import hashlib, json, os, sys
import numpy as np
sys.path.insert(0, os.environ["BAND_ROOT"])
from band_b200 import shard
from oracle import oracle as O
rank, local, world = shard.env_rank()
dist = shard.init_dist("gloo")
N, W, H = int(os.environ["N_FRAMES"]), 64, 48
rng = np.random.default_rng(123)
frames = rng.integers(0, 256, size=(N, O.frame_bytes(W, H, O.FMT_NV12)), dtype=np.uint8)
b, e = shard.shard_slice(N, world, rank)
ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, 22, 22), (O.OP_NORMALIZE, 0, 0, 0, 0, 127.5, 127.5)]
digests = {}
for i in range(b, e):
    st, out = O.run_chain(ops, frames[i], W, H, O.FMT_NV12, O.DT_F32, (1, 22, 22, 3))
    assert st == 0
    digests[i] = hashlib.sha256(out.tobytes()).hexdigest()
mix = shard.mixed_target_plan(N, world, rank, 3)
gathered = [None] * world
dist.all_gather_object(gathered, {"rank": rank, "slice": (b, e), "digests": digests, "mix": mix})
t_max = shard.max_over_ranks(dist, 10.0 + rank)
n_sum = shard.sum_over_ranks(dist, e - b)
dist.barrier()
if rank == 0:
    print(json.dumps({"gathered": gathered, "t_max": t_max, "n_sum": n_sum}))
dist.destroy_process_group()
'''


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("n_frames", [10, 7])
def test_two_rank_gloo_sharding(O, n_frames, tmp_path):
    import json
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    port = _free_port()
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), BAND_ROOT=ROOT, N_FRAMES=str(n_frames))
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.PIPE, text=True))
    outs = [p.communicate(timeout=240) for p in procs]
    for p, (o, e) in zip(procs, outs):
        assert p.returncode == 0, e[-2000:]
    res = json.loads(outs[0][0].strip().splitlines()[-1])
    assert res["t_max"] == 11.0 and res["n_sum"] == n_frames
    # slices are contiguous, disjoint, cover the batch
    slices = sorted(tuple(g["slice"]) for g in res["gathered"])
    assert slices[0][0] == 0 and slices[-1][1] == n_frames and slices[0][1] == slices[1][0]
    # union of shards == single-process result
    rng = np.random.default_rng(123)
    frames = rng.integers(0, 256, size=(n_frames, O.frame_bytes(64, 48, O.FMT_NV12)), dtype=np.uint8)
    ops = [(O.OP_CSC, O.FMT_RGB), (O.OP_RESIZE, 22, 22), (O.OP_NORMALIZE, 0, 0, 0, 0, 127.5, 127.5)]
    merged = {}
    for g in res["gathered"]:
        merged.update({int(k): v for k, v in g["digests"].items()})
    assert sorted(merged) == list(range(n_frames))
    for i in range(n_frames):
        st, out = O.run_chain(ops, frames[i], 64, 48, O.FMT_NV12, O.DT_F32, (1, 22, 22, 3))
        assert hashlib.sha256(out.tobytes()).hexdigest() == merged[i]
    # config-4 dealing: every frame exactly once, target = i % 3, per-rank counts per target differ by <= 1
    seen = {}
    for g in res["gathered"]:
        for t, members in g["mix"].items():
            for i in members:
                assert i % 3 == int(t) and i not in seen
                seen[i] = g["rank"]
    assert sorted(seen) == list(range(n_frames))


def test_shard_slice_properties():
    from band_b200 import shard
    for n in (0, 1, 7, 8, 4096, 4097):
        for world in (1, 2, 4, 8):
            cover = []
            for r in range(world):
                b, e = shard.shard_slice(n, world, r)
                assert 0 <= b <= e <= n
                cover += list(range(b, e))
            assert cover == list(range(n))
    with pytest.raises(ValueError):
        shard.shard_slice(4, 2, 2)
