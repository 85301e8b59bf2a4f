"""Multi-GPU plumbing: scenarios shard by contiguous batch ranges (no data-path collective inside the solve);
the one exchange step is an all-gather of the optimal trajectories.  `torch.distributed`
over NCCL/NVLink on the GPU box, gloo on CPU for the tests."""
import os


def shard_range(total, rank, world):
    """Contiguous range [lo, hi) of rank `rank` of `world`; remainders go to the first ranks."""
    base, rem = divmod(int(total), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def env_world():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def init_process_group(backend=None):
    """Initialise torch.distributed from the torchrun environment (no-op for world size 1)."""
    import torch
    import torch.distributed as dist
    rank, local_rank, world = env_world()
    if world == 1:
        return rank, local_rank, world
    if not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if backend == "nccl":
            torch.cuda.set_device(local_rank)
            dist.init_process_group(backend, rank=rank, world_size=world, device_id=torch.device("cuda", local_rank))
        else:
            dist.init_process_group(backend, rank=rank, world_size=world)
    return rank, local_rank, world


def gather_trajectories(local, counts=None):
    """All-gather of per-rank result rows (U*, optionally X*, J) along dim 0.  Equal shard sizes use one
    all_gather_into_tensor (NCCL over NVLink on the GPU box); ragged shards fall back to all_gather."""
    import torch
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return local
    world = dist.get_world_size()
    if counts is None or len(set(counts)) == 1:
        out = torch.empty((world * local.shape[0],) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(out, local.contiguous())
        return out
    parts = [torch.empty((c,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device) for c in counts]
    dist.all_gather(parts, local.contiguous())
    return torch.cat(parts, dim=0)
