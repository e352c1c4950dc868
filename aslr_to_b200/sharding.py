"""Multi-GPU layout of a batch of independent OCPs.

OCPs never exchange data while they are solved: rank r of G owns the contiguous block
[lo_r, hi_r) of the batch and runs the same kernels on it.  The only collective of the whole path is
the final gather of costs / iteration counts / status and trajectories (`gather_results`), NCCL on
the GPUs and gloo in the CPU-tier tests.  One process per GPU (`torchrun`)."""
import numpy as np


def shard_bounds(B, world, rank):
    """Contiguous block of rank `rank`: the first B % world ranks get one OCP more."""
    if not 0 <= rank < world:
        raise ValueError("rank %d outside world of %d" % (rank, world))
    base, extra = divmod(int(B), int(world))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_sizes(B, world):
    return [shard_bounds(B, world, r)[1] - shard_bounds(B, world, r)[0] for r in range(world)]


def shard_workload(w, world, rank):
    """Slice the per-OCP arrays of a workload dict (x0, K_ocp, mref_ocp); everything else is replicated."""
    x0 = np.atleast_2d(np.asarray(w["x0"], float))
    lo, hi = shard_bounds(x0.shape[0], world, rank)
    out = dict(w)
    out["x0"] = x0[lo:hi]
    for k in ("K_ocp", "mref_ocp"):
        if w.get(k) is not None:
            out[k] = np.asarray(w[k])[lo:hi]
    out["shard"] = (lo, hi)
    return out


def gather_results(local, B, group=None):
    """All-gather per-OCP result tensors.  `local`: dict name -> tensor whose first dimension is this
    rank's shard; returns dict name -> tensor of first dimension B, OCPs in batch order, on every rank.
    Ragged shards (B % world != 0) are padded to the largest shard for the collective."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return dict(local)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    sizes = shard_sizes(B, world)
    nmax = max(sizes)
    out = {}
    for name, t in local.items():
        if t.shape[0] != sizes[rank]:
            raise ValueError("%s: first dimension %d is not this rank's shard size %d" % (name, t.shape[0], sizes[rank]))
        t = t.contiguous()
        if t.shape[0] != nmax:
            pad = torch.zeros((nmax - t.shape[0],) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
            t = torch.cat([t, pad], 0)
        full = torch.empty((world * nmax,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        dist.all_gather_into_tensor(full, t, group=group)
        if all(s == nmax for s in sizes):
            out[name] = full
        else:
            out[name] = torch.cat([full[r * nmax:r * nmax + sizes[r]] for r in range(world)], 0)
    return out
