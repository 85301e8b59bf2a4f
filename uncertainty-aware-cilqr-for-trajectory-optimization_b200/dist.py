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
    all_gather_into_tensor (NCCL over NVLink on the GPU box); ragged shards fall back to all_gather.  `counts` = rows per
    rank; when it is not given the ranks exchange their row counts first (shard_range gives ragged shards whenever the batch is
    not a multiple of the world size)."""
    import torch
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return local
    world = dist.get_world_size()
    if counts is None:
        n = torch.tensor([int(local.shape[0])], dtype=torch.int64, device=local.device)
        ns = [torch.zeros_like(n) for _ in range(world)]
        dist.all_gather(ns, n)
        counts = [int(x.item()) for x in ns]
    if len(set(counts)) == 1:
        out = torch.empty((world * local.shape[0],) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(out, local.contiguous())
        return out
    # ragged shards: every rank pads its rows to the largest shard, one equal-size all-gather, the padding is cut away
    # (all_gather with tensors of different sizes is not supported by every backend)
    cmax = max(counts)
    padded = torch.zeros((cmax,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    padded[:local.shape[0]] = local
    out = torch.empty((world * cmax,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, padded)
    return torch.cat([out[r * cmax:r * cmax + c] for r, c in enumerate(counts)], dim=0)


class _DevArray:
    """A raw device pointer as a CUDA-array-interface object (torch.as_tensor wraps it without copying)."""

    def __init__(self, ptr, shape, typestr="<f8"):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


class PeerGather:
    """The exchange step on the copy engines (cilqr_gather_copy): every rank's [rows, cols] float64 shard into a
    [world * rows, cols] buffer on every rank, by peer-to-peer copies over NVLink instead of an NCCL kernel.  The buffers are
    allocated by the library and opened across the ranks over CUDA IPC; a one-element all-reduce enqueued after the copies is
    the stream-ordered barrier after which a rank's buffer is complete."""

    def __init__(self, solver, rows, cols, device):
        import ctypes as C
        import torch
        import torch.distributed as dist
        from .solver import load_library, _check
        self.lib, self._check, self.solver = load_library(), _check, solver
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.rows, self.cols, self.device = int(rows), int(cols), device
        nbytes = self.world * self.rows * self.cols * 8
        self.own = C.c_void_p()
        handle = (C.c_ubyte * 64)()
        _check(self.lib.cilqr_gather_buffer_create(C.c_int64(nbytes), C.byref(self.own), handle))
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(handle))
        self.ptrs = (C.c_void_p * self.world)()
        self.opened = []
        for r, hb in enumerate(handles):
            if r == self.rank:
                self.ptrs[r] = self.own.value
            else:
                q = C.c_void_p()
                _check(self.lib.cilqr_gather_buffer_open((C.c_ubyte * 64).from_buffer_copy(hb), C.byref(q)))
                self.ptrs[r] = q.value
                self.opened.append(q)
        self.gathered = torch.as_tensor(_DevArray(self.own.value, (self.world * self.rows, self.cols)), device=device)
        self.token = torch.zeros(1, dtype=torch.float32, device=device)
        dist.barrier()

    def gather(self, local):
        """Enqueue on torch's current stream: this rank's shard to every rank, then the barrier."""
        import ctypes as C
        import torch
        import torch.distributed as dist
        assert local.is_contiguous() and tuple(local.shape) == (self.rows, self.cols)
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        self._check(self.lib.cilqr_gather_copy(self.solver._h, C.c_int(self.world), C.c_int(self.rank), C.c_void_p(local.data_ptr()),
                                               C.c_int64(self.rows * self.cols), self.ptrs, st))
        dist.all_reduce(self.token)  # one element: every rank's copies precede it on that rank's stream
        return self.gathered

    def close(self):
        import torch
        import torch.distributed as dist
        torch.cuda.synchronize()
        dist.barrier()  # nobody unmaps while a peer still writes
        for q in self.opened:
            self.lib.cilqr_gather_buffer_close(q)
        self.opened = []
        dist.barrier()
        if self.own:
            self.lib.cilqr_gather_buffer_destroy(self.own)
            self.own = None
