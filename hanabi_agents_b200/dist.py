"""The little that is distributed: one process per GPU, games / replay shards partitioned by
rank, one all-reduce of the gradient bucket per learner step.  torch.distributed is
plumbing only (NCCL over NVLink on the B200 box, gloo in the CPU tests)."""
import os


def env_world():
    """(rank, local_rank, world) from the torchrun environment (1 process = 1 GPU)."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")),
            int(os.environ.get("WORLD_SIZE", "1")))


def shard(n_total, rank, world):
    """Contiguous shard [base, base + n) of n_total independent units (games, replay rows) for
    `rank`; the first n_total % world ranks hold one unit more."""
    if not (0 <= rank < world) or n_total < 0:
        raise ValueError("bad shard request")
    q, r = divmod(n_total, world)
    n = q + (1 if rank < r else 0)
    base = rank * q + min(rank, r)
    return base, n


def allreduce_sum_(tensor, group=None):
    """In-place sum over ranks (the learner's only collective); no-op without a process group."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(tensor, op=dist.ReduceOp.SUM, group=group)
    return tensor


def max_over_ranks(value, device="cpu", group=None):
    """max over ranks of a python float (timings are reported as the slowest rank's)."""
    import torch
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        t = torch.tensor([float(value)], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
        return float(t.item())
    return float(value)


def sum_over_ranks(value, device="cpu", group=None):
    import torch
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        t = torch.tensor([float(value)], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
        return float(t.item())
    return float(value)


class NvlinkAllReduce:
    """The gradient all-reduce as ONE kernel over NVLink peer memory (csrc/comm.cu): every rank maps
    its peers' buckets with CUDA IPC; `__call__` sums the bucket in place, identically on every rank.
    torch.distributed only carries the 128-byte handles once.  Raises if the peers cannot be mapped
    (ranks on different nodes, no peer access): the caller then stays on NCCL."""

    def __init__(self, bucket, group=None):
        import ctypes as C
        import torch
        import torch.distributed as dist
        from . import _lib
        if not (dist.is_available() and dist.is_initialized()):
            raise RuntimeError("NvlinkAllReduce needs an initialised process group")
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self._bucket = bucket
        self._h = C.c_void_p()
        L = _lib.lib()
        _lib.check(L.hb_comm_create(self.rank, self.world, bucket.device.index, bucket.data_ptr(), bucket.numel(),
                                    C.byref(self._h)))
        mine = (C.c_char * 128)()
        _lib.check(L.hb_comm_export(self._h, mine))
        gathered = [None] * self.world
        dist.all_gather_object(gathered, bytes(mine.raw), group=group)
        blob = b"".join(gathered)
        ok = 1
        try:
            _lib.check(L.hb_comm_open(self._h, blob))
        except Exception as e:      # noqa: BLE001 -- every rank must learn about it
            ok, self._err = 0, e
        flag = torch.tensor([ok], device=bucket.device, dtype=torch.int32)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        if int(flag.item()) == 0:
            L.hb_comm_destroy(self._h)
            self._h = None
            raise RuntimeError("peer buckets cannot be mapped: %r" % (getattr(self, "_err", "a peer failed"),))

    def __call__(self, stream=None):
        from . import _lib
        _lib.check(_lib.lib().hb_comm_allreduce(self._h, _lib.current_stream_ptr() if stream is None else stream))

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            try:
                from . import _lib
                _lib.lib().hb_comm_destroy(h)
            except Exception:
                pass
            self._h = None
