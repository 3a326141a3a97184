"""The N>1 host logic on CPU: two gloo ranks shard the A-step pair list, all-gather the coordinate
slabs and the rows, and must reproduce the single-process result bit for bit (the numpy oracle
stands in for the GPU kernel here; the GPU tests cover the kernel itself)."""
import os
import socket
import sys

import numpy as np

from conftest import ROOT
from igm_b200 import parallel, synth


def test_structure_shard_covers_everything():
    for n, ws in [(1000, 8), (10, 3), (7, 8), (1250, 1)]:
        ranges = [parallel.structure_shard(n, r, ws) for r in range(ws)]
        assert ranges[0][0] == 0 and ranges[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(ranges, ranges[1:]))
        sizes = [hi - lo for lo, hi in ranges]
        assert max(sizes) - min(sizes) <= 1


def test_pair_shards_balance_cost_and_stay_contiguous():
    rng = np.random.default_rng(0)
    cost = rng.choice([1, 2, 4], 10001)
    for ws in (1, 2, 3, 8):
        sh = parallel.pair_shards(cost, ws)
        assert sh[0][0] == 0 and sh[-1][1] == cost.size
        assert all(a[1] == b[0] for a, b in zip(sh, sh[1:]))
        tot = [cost[lo:hi].sum() for lo, hi in sh]
        assert max(tot) - min(tot) <= 8
    assert parallel.pair_shards([], 4) == [(0, 0)] * 4


def _worker(rank, ws, port, out_dir):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    from oracle import astep_oracle as ao
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    g = synth.Genome([9, 7, 6, 4, 3], n_autosomes=3)
    rad = synth.bead_radii(len(g), 1500.0, 0.3)
    S_total = 25                                   # does not divide by the world size: blocks of 13 and 12 structures
    lo, hi = parallel.structure_shard(S_total, rank, ws)
    full = synth.population_territories(g, S_total, R=1500.0, seed=3)
    local = torch.from_numpy(np.ascontiguousarray(full[:, lo:hi, :]))
    slab, offs = parallel.all_gather_slab(local, return_offsets=True)
    slab = slab.numpy()
    assert np.array_equal(slab, full) and offs == [0, 13, 25]
    pi, pj, pw = synth.hic_pairs(g, 0.004, seed=1)
    pl = np.zeros(pi.size)

    def fn(a, b, c, d):
        return ao.actdist_rows(a, b, c, d, slab, rad, g.chrom, g.copy_index, 2.0)

    rows = parallel.distributed_actdist(fn, pi, pj, pw, pl, g.chrom, g.copy_off)
    np.savez(os.path.join(out_dir, "rank%d.npz" % rank), row=rows[0], col=rows[1], dist=rows[2], prob=rows[3])
    # the structure-sharded variant: nobody holds the whole population, counts and radix histograms are all-reduced
    from oracle.astep_shard_oracle import NumpyShard
    shard = NumpyShard(np.ascontiguousarray(full[:, lo:hi, :]), rad, g.chrom, g.copy_index)
    rows2 = parallel.sharded_actdist(shard, S_total, pi, pj, pw, pl, 2.0, pair_chunk=37)
    np.savez(os.path.join(out_dir, "shard%d.npz" % rank), row=rows2[0], col=rows2[1], dist=rows2[2], prob=rows2[3])
    dist.destroy_process_group()


def test_two_rank_astep_equals_single_process(tmp_path):
    import torch.multiprocessing as mp
    from oracle import astep_oracle as ao
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    g = synth.Genome([9, 7, 6, 4, 3], n_autosomes=3)
    rad = synth.bead_radii(len(g), 1500.0, 0.3)
    full = synth.population_territories(g, 25, R=1500.0, seed=3)
    pi, pj, pw = synth.hic_pairs(g, 0.004, seed=1)
    want = ao.actdist_rows(pi, pj, pw, np.zeros(pi.size), full, rad, g.chrom, g.copy_index, 2.0)
    assert want[0].size > 50
    for r in range(2):
        for variant in ("rank", "shard"):          # slab all-gather + pair sharding; structure sharding + all-reduced histograms
            got = np.load(os.path.join(str(tmp_path), "%s%d.npz" % (variant, r)))
            for k, name in enumerate(("row", "col", "dist", "prob")):
                assert np.array_equal(got[name].view(np.uint32), want[k].view(np.uint32)), (r, variant, name)
