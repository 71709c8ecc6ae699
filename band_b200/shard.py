"""Frame sharding across the GPUs of one box: frames are independent, so rank r takes a
contiguous slice of the batch and runs it on its own GPU with its own processor and stream. There is NO data-path
collective; torch.distributed carries only the barrier and the max-over-ranks of the measured time."""
import os


def shard_slice(n_frames, world, rank):
    """[begin, end) of rank's contiguous slice: frame i goes to rank i // ceil(n/world)."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError("bad world/rank")
    per = -(-n_frames // world)
    begin = min(n_frames, rank * per)
    return begin, min(n_frames, begin + per)


def mixed_target_plan(n_frames, world, rank, n_targets):
    """BASELINE config 4: frame i uses target i % n_targets. Frames are dealt so that every rank gets the same
    target mix: rank r takes frames r, r + world, ... of EACH target class. Returns {target: [frame indices]}."""
    plan = {t: [] for t in range(n_targets)}
    for t in range(n_targets):
        members = list(range(t, n_frames, n_targets))
        plan[t] = members[rank::world]
    return plan


def env_rank():
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")),
            int(os.environ.get("WORLD_SIZE", "1")))


def init_dist(backend):
    """Process group for world > 1 (nccl on the GPU box, gloo in CPU tests); None for a single process."""
    rank, local, world = env_rank()
    if world <= 1:
        return None
    import torch.distributed as dist
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29511")
    if backend == "nccl":
        import torch
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        dist.init_process_group(backend)
    return dist


def max_over_ranks(dist, value, device="cpu"):
    if dist is None:
        return float(value)
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(dist, value, device="cpu"):
    if dist is None:
        return float(value)
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())
