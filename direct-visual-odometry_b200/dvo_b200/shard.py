"""Multi-GPU plumbing for batched independent sequences: static sharding of sequence ids
across ranks (id mod world), one process + one dvo_ctx per GPU, NO data-path collective — sequences never
exchange pixels.  torch.distributed is used only for the bench's barrier / max-over-ranks timing and for
gathering per-sequence results on rank 0 (NCCL on GPUs, gloo in the CPU tests)."""
import numpy as np


def shard_ids(n_sequences, world, rank):
    """Sequence ids owned by `rank`: id mod world == rank."""
    if not (0 <= rank < world):
        raise ValueError("rank %d outside world %d" % (rank, world))
    return np.arange(rank, n_sequences, world, dtype=np.int64)


def owner_of(seq_id, world):
    return int(seq_id) % int(world)


def max_over_ranks(value, dist=None, device="cpu"):
    """Device-timed milliseconds -> the slowest rank's (what a whole-job throughput must be divided by)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return float(value)
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def whole_job_rate(units_per_rank, ms, dist=None, device="cpu"):
    """Aggregate throughput: units processed by all ranks / slowest rank's time (units per second)."""
    total = float(units_per_rank)
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        import torch
        t = torch.tensor([total], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        total = float(t.item())
    return total / (max_over_ranks(ms, dist, device) * 1e-3)


def gather_by_sequence(local_ids, local_values, n_sequences, dist=None):
    """Per-sequence results (e.g. final poses [n_local, ...]) of every rank reassembled in sequence-id order on
    every rank.  Host-side, after the timed region."""
    local_values = np.asarray(local_values)
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        out = np.empty((n_sequences,) + local_values.shape[1:], local_values.dtype)
        out[np.asarray(local_ids)] = local_values
        return out
    parts = [None] * dist.get_world_size()
    dist.all_gather_object(parts, (np.asarray(local_ids), local_values))
    out = np.empty((n_sequences,) + local_values.shape[1:], local_values.dtype)
    seen = np.zeros(n_sequences, bool)
    for ids, vals in parts:
        if seen[ids].any():
            raise RuntimeError("a sequence id is owned by two ranks")
        out[ids] = vals; seen[ids] = True
    if not seen.all():
        raise RuntimeError("sequence ids %s are owned by no rank" % np.flatnonzero(~seen)[:8].tolist())
    return out
