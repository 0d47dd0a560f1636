"""Data-parallel plumbing for the op: the batch shards naturally (no output element
depends on another batch item, ms_deform_im2col_cuda.cuh:256-270), so ranks take disjoint
batch slices and there is NO data-path collective.  torch.distributed is used only for the
barrier around a timed region and for the max-over-ranks of a duration (the reference's
only distributed machinery is DDP around the whole model, main.py:82-88)."""
import torch
import torch.distributed as dist


def shard_range(n_items, rank, world):
    """Contiguous, balanced [start, stop) of a batch of n_items for `rank` of `world`
    (the first n_items % world ranks get one extra item; empty shards are allowed)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world %r/%r" % (rank, world))
    base, extra = divmod(int(n_items), world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def shard_batch(tensors, rank, world):
    """Slice every tensor whose leading dim is the batch; index tensors pass through."""
    n = tensors["value"].shape[0]
    a, b = shard_range(n, rank, world)
    return {k: (v[a:b].contiguous() if k in ("value", "loc", "aw", "grad_out") else v)
            for k, v in tensors.items()}


def max_over_ranks(x, device=None):
    """max of a python float over all ranks (identity when not distributed)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(x)
    t = torch.tensor([float(x)], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(x, device=None):
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(x)
    t = torch.tensor([float(x)], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())
