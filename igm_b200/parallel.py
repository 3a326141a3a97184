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


def all_gather_slab(local, group=None, return_offsets=False):
    """local: torch tensor [n_bead][S_local][3]; S_local may differ between ranks (``structure_shard`` gives blocks that
    differ by one when the population does not divide evenly).  -> [n_bead][sum S_local][3], the structures of rank r
    at [off[r], off[r+1]); with ``return_offsets`` also the list ``off`` (world + 1 entries)."""
    import torch
    d = _dist()
    if d is None or d.get_world_size(group) == 1:
        return (local, [0, int(local.shape[1])]) if return_offsets else local
    ws = d.get_world_size(group)
    n = torch.tensor([local.shape[1]], dtype=torch.int64, device=local.device)
    sizes = [torch.zeros_like(n) for _ in range(ws)]
    d.all_gather(sizes, n, group=group)
    sizes = [int(v.item()) for v in sizes]
    cap = max(sizes)
    if min(sizes) == cap:
        parts = [torch.empty_like(local) for _ in range(ws)]
        d.all_gather(parts, local.contiguous(), group=group)
    else:                                   # pad every block to the largest, gather, cut the padding off again
        padded = torch.zeros((local.shape[0], cap, local.shape[2]), dtype=local.dtype, device=local.device)
        padded[:, :local.shape[1], :] = local
        parts = [torch.empty_like(padded) for _ in range(ws)]
        d.all_gather(parts, padded, group=group)
        parts = [p[:, :sizes[r], :] for r, p in enumerate(parts)]
    slab = torch.cat(parts, dim=1).contiguous()
    off = [0]
    for v in sizes:
        off.append(off[-1] + v)
    return (slab, off) if return_offsets else slab


def total_structures(n_local, device=None, group=None):
    """sum of the per-rank structure counts"""
    import torch
    d = _dist()
    if d is None or d.get_world_size(group) == 1:
        return int(n_local)
    t = torch.tensor([int(n_local)], dtype=torch.int64, device=device)
    d.all_reduce(t, op=d.ReduceOp.SUM, group=group)
    return int(t.item())


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


class GpuShard(object):
    """One structure shard on a GPU for ``sharded_actdist``: a Population holding this shard's structures plus the
    device buffers the phases exchange."""

    def __init__(self, pop, cap):
        import torch
        self.pop = pop
        dev = torch.device("cuda:%d" % pop.device)
        # uint32 payloads in int32 tensors: the sums stay below 2^31 (at most 16 copy pairs x S_total per pair)
        self.counts = torch.zeros(cap, dtype=torch.int32, device=dev)
        self.hist = torch.zeros((cap, 256), dtype=torch.int32, device=dev)
        self.want = torch.zeros(cap, dtype=torch.int32, device=dev)
        self.prefix = torch.zeros(cap, dtype=torch.int32, device=dev)
        self.kernel_ms = 0.0
        self.m = 0

    def set_pairs(self, pi, pj, pw, pl):
        self.m = int(pi.size)
        self.pop.set_pairs(pi, pj, pw, pl)

    def count(self, contact_range):
        self.pop.shard_count(contact_range, self.counts.data_ptr())
        self.kernel_ms += self.pop.last_kernel_ms()
        return self.counts[:self.m]

    def rank(self, n_struct_total, total_counts):
        self.pop.shard_rank(n_struct_total, total_counts.data_ptr(), self.want.data_ptr(), self.prefix.data_ptr())

    def histogram(self, radix_pass):
        self.pop.shard_hist(radix_pass, self.prefix.data_ptr(), self.hist.data_ptr())
        if radix_pass == 3:
            self.kernel_ms += self.pop.last_kernel_ms()
        return self.hist[:self.m]

    def pick(self, radix_pass, total_hist):
        self.pop.shard_pick(radix_pass, total_hist.data_ptr(), self.want.data_ptr(), self.prefix.data_ptr())

    def sync(self):
        import torch
        torch.cuda.synchronize(self.counts.device)

    def collect(self):
        return self.pop.actdist_collect()


def sharded_actdist(shards, n_struct_total, pair_i, pair_j, pwish, plast, contact_range=2.0, group=None, pair_chunk=1 << 17):
    """Hi-C activation distances with the population SHARDED by structure (BASELINE configs[4]; north_star: "NCCL ...
    reduces the per-pair distance histograms").  ``shards``: Population(s) of this process, each holding only its own
    structures -- one per rank in production; several on one GPU emulate several ranks (their buffers are added before
    the all-reduce).  (Objects with the methods of ``GpuShard`` are accepted as they are: the CPU tests pass numpy
    stand-ins under gloo.)  Per chunk of ``pair_chunk`` pairs: local contact counts -> all-reduce(sum) -> probability and
    wanted rank o -> four radix-256 passes, each a local histogram [pairs][256] -> all-reduce(sum) -> next digit of the
    o-th smallest squared distance.  Every rank ends with the full row list, bit-identical to ``Population.actdist``
    over the whole population.  -> (row, col, dist, prob)"""
    d = _dist()
    shards = list(shards) if isinstance(shards, (list, tuple)) else [shards]
    pair_i = np.ascontiguousarray(pair_i, dtype=np.int32)
    pair_j = np.ascontiguousarray(pair_j, dtype=np.int32)
    pwish = np.ascontiguousarray(pwish, dtype=np.float64)
    plast = np.ascontiguousarray(plast, dtype=np.float64)
    n = int(min(pair_chunk, max(pair_i.size, 1)))
    shards = [sh if hasattr(sh, "histogram") else GpuShard(sh, n) for sh in shards]

    def total(parts):
        t = parts[0]
        for b in parts[1:]:
            t += b
        if d is not None and d.get_world_size(group) > 1:
            d.all_reduce(t, op=d.ReduceOp.SUM, group=group)
        for sh in shards:
            sh.sync()
        return t

    out = [[], [], [], []]
    for lo in range(0, pair_i.size, n):
        hi = min(pair_i.size, lo + n)
        for sh in shards:
            sh.set_pairs(pair_i[lo:hi], pair_j[lo:hi], pwish[lo:hi], plast[lo:hi])
        tot = total([sh.count(contact_range) for sh in shards])
        for sh in shards:
            sh.rank(n_struct_total, tot)
        for radix_pass in range(4):
            tot = total([sh.histogram(radix_pass) for sh in shards])
            for sh in shards:
                sh.pick(radix_pass, tot)
        for k, a in enumerate(shards[0].collect()):
            out[k].append(np.array(a, copy=True))
    sharded_actdist.last_kernel_ms = sum(getattr(sh, "kernel_ms", 0.0) for sh in shards)
    if not out[0]:
        return (np.zeros(0, np.int32), np.zeros(0, np.int32), np.zeros(0, np.float32), np.zeros(0, np.float32))
    return tuple(np.concatenate(o) for o in out)
