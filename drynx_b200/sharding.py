"""Multi-GPU plumbing: data providers (and proofs) are independent, so each rank owns a contiguous block and the
only exchange is a gather of the per-rank partial aggregate (V x 128 bytes), folded on rank 0 with the same
aggregation entry point. torch.distributed is transport only (NCCL on GPUs, gloo in the CPU tests)."""
import numpy as np


def shard_range(n, rank, world):
    """Contiguous block partition of n units: rank r owns [lo, hi); sizes differ by at most one."""
    base, extra = divmod(n, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def gather_partials(partial, world, dist, device=None):
    """All-gathers one serialised CipherVector (uint8[V*128]) per rank; returns uint8[world][V*128] on every rank."""
    import torch
    t = torch.as_tensor(np.ascontiguousarray(partial, dtype=np.uint8))
    if device is not None:
        t = t.to(device)
    out = torch.empty(world * t.numel(), dtype=torch.uint8, device=t.device)
    dist.all_gather_into_tensor(out, t)
    return out.view(world, -1)


def fold_on_root(gathered, rank, aggregate):
    """Rank 0 folds the gathered partials with `aggregate` (Backend.aggregate or any CipherVector.Add stand-in)."""
    if rank != 0:
        return None
    rows = gathered.cpu().numpy() if hasattr(gathered, 'cpu') else np.asarray(gathered)
    return aggregate([rows[r] for r in range(rows.shape[0])])
