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
