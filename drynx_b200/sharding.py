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


def gather_and_fold(partial, rank, world, dist, fold, gathered=None):
    """The one exchange step of a sharded survey. `partial` is this rank's partial aggregate as a uint8 torch tensor
    [V * ct_bytes] (a CUDA tensor under NCCL, a CPU tensor under gloo); every rank contributes it to an all-gather and
    rank 0 folds the `world` partials with `fold(gathered.view(world, -1))` (drx_aggregate_dev on the GPUs, any
    CipherVector.Add stand-in in the CPU tests) and returns the result; the other ranks return None. With one rank
    there is nothing to gather and the partial is the total. `gathered` may supply the receive buffer."""
    import torch
    if world == 1:
        return partial
    if gathered is None:
        gathered = torch.empty(world * partial.numel(), dtype=torch.uint8, device=partial.device)
    dist.all_gather_into_tensor(gathered, partial)
    if rank != 0:
        return None
    return fold(gathered.view(world, -1))


def gather_accept_bytes(ok, n_total, rank, world, dist):
    """Verification results of a batch split by proof index with shard_range: `ok` holds this rank's accept bytes
    (uint8 torch tensor); returns the n_total accept bytes in proof order on every rank (shards differ in length by at
    most one, so they travel padded to the longest)."""
    import torch
    if world == 1:
        return ok
    longest = (n_total + world - 1) // world
    buf = torch.zeros(longest, dtype=torch.uint8, device=ok.device)
    buf[:ok.numel()] = ok
    out = torch.empty(world * longest, dtype=torch.uint8, device=ok.device)
    dist.all_gather_into_tensor(out, buf)
    parts = []
    for r in range(world):
        lo, hi = shard_range(n_total, r, world)
        parts.append(out[r * longest:r * longest + (hi - lo)])
    return torch.cat(parts)
