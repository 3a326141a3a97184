"""Sharding of the A/M iteration over the GPUs of one box (one process per GPU).

The reference parallelises by mapping independent tasks over workers and exchanging results
through the file system (igm/core/step.py:259-274, igm/steps/ModelingStep.py:97,
ActivationDistanceStep.py:189).  Here:

* M-step: structures are split by index, contiguously; no data-path collective.
* A-step: every haploid pair needs all structures, so the per-rank coordinate slabs
  ``[n_bead][S_local][3]`` are all-gathered on the structure axis (NCCL over NVLink on GPUs, gloo
  in the CPU tests), the pair list is split into contiguous chunks of equal estimated cost, and the
  actdist rows of the chunks are all-gathered; concatenating them in rank order reproduces the
  single-process row order exactly.

Everything takes an optional ``group``; with ``torch.distributed`` uninitialised (world size 1) the
functions degrade to the identity, so single-GPU callers use the same code path.
"""
import numpy as np


def _dist():
    import torch.distributed as dist
    return dist if dist.is_available() and dist.is_initialized() else None


def world(group=None):
    d = _dist()
    if d is None:
        return 0, 1
    return d.get_rank(group), d.get_world_size(group)


def structure_shard(n_struct, rank, world_size):
    """Contiguous range of structure indices of ``rank`` (sizes differ by at most one)."""
    base, rem = divmod(int(n_struct), int(world_size))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def pair_shards(cost, world_size):
    """Split pairs 0..n-1 into ``world_size`` contiguous chunks of (nearly) equal total ``cost``.
    -> list of (lo, hi)."""
    cost = np.asarray(cost, dtype=np.float64)
    n = cost.size
    if n == 0:
        return [(0, 0)] * world_size
    csum = np.cumsum(cost)
    cuts = [0]
    for r in range(1, world_size):
        cuts.append(int(np.searchsorted(csum, csum[-1] * r / world_size, side="left")))
    cuts.append(n)
    cuts = np.maximum.accumulate(cuts)
    return [(int(cuts[r]), int(cuts[r + 1])) for r in range(world_size)]


def pair_cost(pair_i, pair_j, chrom, copy_off):
    """Work of one haploid pair = number of copy combinations get_actdist evaluates
    (ActivationDistanceStep.py:382-413): zip for intra, product for inter."""
    ni = np.diff(copy_off)[pair_i]
    nj = np.diff(copy_off)[pair_j]
    intra = np.asarray(chrom)[pair_i] == np.asarray(chrom)[pair_j]
    return np.where(intra, np.minimum(ni, nj), ni * nj)


def all_gather_slab(local, group=None):
    """local: torch tensor [n_bead][S_local][3] (same S_local on every rank) ->
    [n_bead][S_local * world][3], structures of rank r at [r * S_local, (r+1) * S_local)."""
    import torch
    d = _dist()
    if d is None or d.get_world_size(group) == 1:
        return local
    parts = [torch.empty_like(local) for _ in range(d.get_world_size(group))]
    d.all_gather(parts, local.contiguous(), group=group)
    return torch.cat(parts, dim=1).contiguous()


def all_gather_rows(rows, device=None, group=None):
    """rows: tuple of equally long 1-D numpy arrays (this rank's chunk).  -> tuple of the
    concatenation over ranks in rank order.  Works with gloo (CPU tensors) and NCCL (pass the
    rank's CUDA device)."""
    import torch
    d = _dist()
    if d is None or d.get_world_size(group) == 1:
        return rows
    ws = d.get_world_size(group)
    n = torch.tensor([rows[0].size], dtype=torch.int64, device=device)
    sizes = [torch.zeros_like(n) for _ in range(ws)]
    d.all_gather(sizes, n, group=group)
    sizes = [int(s.item()) for s in sizes]
    cap = max(max(sizes), 1)
    out = []
    for a in rows:
        a = np.ascontiguousarray(a)
        # ship the raw bytes so that float32 bit patterns survive unchanged
        buf = torch.zeros(cap * a.dtype.itemsize, dtype=torch.uint8, device=device)
        if a.size:
            buf[:a.nbytes] = torch.from_numpy(a.view(np.uint8).copy()).to(buf.device)
        parts = [torch.empty_like(buf) for _ in range(ws)]
        d.all_gather(parts, buf, group=group)
        out.append(np.concatenate([parts[r][:sizes[r] * a.dtype.itemsize].cpu().numpy().view(a.dtype) for r in range(ws)]))
    return tuple(out)


def distributed_actdist(actdist_fn, pair_i, pair_j, pwish, plast, chrom, copy_off, device=None, group=None):
    """Run ``actdist_fn(pair_i, pair_j, pwish, plast) -> (row, col, dist, prob)`` on this rank's
    chunk of the pair list (the full population must be visible to it) and all-gather the rows.
    The result is identical, row for row, to ``actdist_fn`` over the whole list."""
    rank, ws = world(group)
    pair_i = np.asarray(pair_i)
    pair_j = np.asarray(pair_j)
    lo, hi = pair_shards(pair_cost(pair_i, pair_j, chrom, copy_off), ws)[rank]
    rows = actdist_fn(pair_i[lo:hi], pair_j[lo:hi], np.asarray(pwish)[lo:hi], np.asarray(plast)[lo:hi])
    return all_gather_rows(tuple(rows), device=device, group=group)
