"""A-step parity (north_star criterion 2: bit-exact): CUDA activation distances and contact
assignment against outputs of the reference's own Python (golden fixture) and the numpy oracle."""
import numpy as np
import pytest

from igm_b200 import synth
from igm_b200.astep import Population, unpack_bits
from oracle import astep_oracle as ao

pytestmark = pytest.mark.gpu


def _pop(g, device=0):
    p = Population(g["coordinates"].shape[1], g["radii"], g["chrom"], g["copy_off"], g["copy_beads"], device=device)
    p.upload(np.ascontiguousarray(g["coordinates"]))
    return p


def _bits_equal(a, b):
    return all(np.array_equal(x.view(np.uint32), y.view(np.uint32)) for x, y in zip(a, b))


def test_actdist_golden_bit_exact(astep_golden):
    g = astep_golden
    with _pop(g) as pop:
        got = pop.actdist(g["pair_i"], g["pair_j"], g["pwish"], g["plast"], 2.0)
    assert _bits_equal(got, (g["row"], g["col"], g["dist"], g["prob"]))


def test_assign_golden_bonds(astep_golden):
    g = astep_golden
    with _pop(g) as pop:
        bits = pop.assign(g["row"], g["col"], g["dist"])
    mask = unpack_bits(bits, g["coordinates"].shape[1])
    assert np.array_equal(mask, ao.assign(g["coordinates"], g["row"], g["col"], g["dist"]))
    for s in g["assign_structs"]:
        got = np.stack([g["row"][mask[:, s]], g["col"][mask[:, s]]], axis=1)
        got = got[np.lexsort((got[:, 1], got[:, 0]))]
        assert np.array_equal(got, g["bonds_%d" % s])


def test_download_roundtrip(astep_golden):
    with _pop(astep_golden) as pop:
        assert np.array_equal(pop.download(), astep_golden["coordinates"])


@pytest.mark.parametrize("n_struct,seed,force_smem", [(1, 1, False), (33, 2, False), (257, 3, False), (1000, 4, False),
                                                      (1500, 5, False), (2100, 6, False), (1000, 7, True), (777, 8, "wide")])
def test_actdist_and_assign_vs_oracle_synthetic(n_struct, seed, force_smem, monkeypatch):
    """ragged structure counts (not multiples of 32 / 256), X/Y single-copy bins, plast in {0, <1, 1};
    covers the register-resident kernel (<= 2048 structures) and the shared-memory one"""
    if force_smem:
        monkeypatch.setenv("IGMK_ACTDIST_SMEM", "1")
    if force_smem == "wide":                          # the 1024-thread form of the shared-memory kernel
        monkeypatch.setenv("IGMK_ACTDIST_NT", "1024")
    g = synth.Genome([9, 7, 6, 4, 3], n_autosomes=3)
    rad = synth.bead_radii(len(g), 1500.0, 0.3)
    crd = synth.population_territories(g, n_struct, R=1500.0, seed=seed)
    pi, pj, pw = synth.hic_pairs(g, 0.004, seed=seed)
    rng = np.random.default_rng(seed)
    pl = np.where(rng.random(pi.size) < 0.4, 0.0, np.round(rng.random(pi.size) * pw, 4))
    pl[::13] = 1.0
    # a few degenerate inputs: i == j is skipped by the reference (ActivationDistanceStep.py:351)
    pi = np.concatenate([pi, [3, 0]]).astype(np.int32)
    pj = np.concatenate([pj, [3, 0]]).astype(np.int32)
    pw = np.concatenate([pw, [0.5, 1.0]])
    pl = np.concatenate([pl, [0.0, 0.0]])
    copy_index = g.copy_index
    want = ao.actdist_rows(pi, pj, pw, pl, crd, rad, g.chrom, copy_index, 2.0)
    with Population(n_struct, rad, g.chrom, g.copy_off, g.copy_beads) as pop:
        pop.upload(crd)
        got = pop.actdist(pi, pj, pw, pl, 2.0)
        assert got[0].size == want[0].size and got[0].size > 0
        assert _bits_equal(got, want)
        bits = pop.assign(*got[:3])
        assert np.array_equal(bits, ao.pack_bits(ao.assign(crd, *got[:3])))


def test_assign_ties_at_the_activation_distance():
    """the row's own defining pair sits within an ulp of dist: exercise exact ties by setting
    dist to the float32 norm itself and to its neighbours"""
    g = synth.Genome([6, 5], n_autosomes=1)
    rad = synth.bead_radii(len(g), 800.0, 0.3)
    S = 64
    crd = synth.population_territories(g, S, R=800.0, seed=9)
    row = np.repeat(np.arange(0, 5, dtype=np.int32), 3)
    col = row + 6
    base = ao.particle_distance(crd[row, 7], crd[col, 7])
    dist = base.copy()
    dist[1::3] = np.nextafter(base[1::3], np.float32(0))
    dist[2::3] = np.nextafter(base[2::3], np.float32(np.inf))
    with Population(S, rad, g.chrom, g.copy_off, g.copy_beads) as pop:
        pop.upload(crd)
        bits = pop.assign(row, col, dist)
    mask = unpack_bits(bits, S)
    assert np.array_equal(mask, ao.assign(crd, row, col, dist))
    assert mask[0::3, 7].all() and not mask[1::3, 7].any() and mask[2::3, 7].all()


def test_empty_batches(astep_golden):
    with _pop(astep_golden) as pop:
        got = pop.actdist([], [], [], [], 2.0)
        assert all(a.size == 0 for a in got)
        assert pop.assign([], [], []).shape[0] == 0


def test_bad_indices_are_rejected(astep_golden):
    with _pop(astep_golden) as pop:
        with pytest.raises(RuntimeError, match="out of range"):
            pop.actdist([0], [10 ** 6], [0.5], [0.0])
        with pytest.raises(RuntimeError, match="out of range"):
            pop.assign([0], [10 ** 6], [1.0])


def test_full_size_population_properties():
    """BASELINE configs[1] size (3008 beads x 1000 structures): properties that need no CPU replay of
    the whole batch -- the emitted distance IS the o-th order statistic of the pair's squared distances
    (checked in numpy on a sample), launches repeat bit for bit, and splitting the pair list (what the
    multi-GPU path does) changes nothing."""
    g = synth.genome_2mb()
    rad = synth.bead_radii(len(g))
    S = 1000
    crd = synth.population_territories(g, S, seed=77)
    pi, pj, pw = synth.hic_pairs(g, 0.02, seed=5)
    pl = np.zeros(pi.size)
    with Population(S, rad, g.chrom, g.copy_off, g.copy_beads) as pop:
        pop.upload(crd)
        a = pop.actdist(pi, pj, pw, pl, 2.0)
        b = pop.actdist(pi, pj, pw, pl, 2.0)
        assert _bits_equal(a, b) and a[0].size > 10000
        cut = pi.size // 3
        parts = [pop.actdist(pi[s], pj[s], pw[s], pl[s], 2.0) for s in (slice(0, cut), slice(cut, None))]
        assert _bits_equal(a, tuple(np.concatenate([p[k] for p in parts]) for k in range(4)))
        bits = pop.assign(a[0], a[1], a[2])
        assert np.array_equal(bits, pop.assign(a[0], a[1], a[2]))
    # order-statistic property on a sample of pairs, from first principles
    rng = np.random.default_rng(0)
    checked = 0
    for q in rng.choice(pi.size, 400, replace=False):
        i, j = int(pi[q]), int(pj[q])
        rows = np.nonzero((a[0] == g.copy_index[i][0]) & (a[1] == g.copy_index[j][0]))[0]
        if rows.size == 0:
            continue                                   # corrected probability was 0: no row emitted
        dist, prob = float(a[2][rows[0]]), float(a[3][rows[0]])
        ii, jj = g.copy_index[i], g.copy_index[j]
        pairs = list(zip(ii, jj)) if g.chrom[i] == g.chrom[j] else [(x, y) for x in ii for y in jj]
        d2 = np.concatenate([((crd[x] - crd[y]) ** 2).sum(axis=1) for x, y in pairs]).astype(np.float64)
        n = d2.size
        # prob and dist went through a 4-decimal text round trip: bracket the rank accordingly
        below = np.count_nonzero(np.sqrt(d2) <= dist - 1e-3)
        upto = np.count_nonzero(np.sqrt(d2) <= dist + 1e-3)
        o_lo, o_hi = int(np.floor(n * (prob - 6e-5))) - 1, int(np.ceil(n * (prob + 6e-5))) + 1
        assert below <= o_hi + 1 and upto >= min(o_lo, n - 1) + 1, (i, j, below, upto, o_lo, o_hi)
        checked += 1
    assert checked > 40
    # assignment bits: a structure is assigned iff its distance is within the activation distance
    mask = unpack_bits(bits[:200], S)
    for r in range(0, 200, 7):
        d = np.sqrt(((crd[a[0][r]].astype(np.float64) - crd[a[1][r]]) ** 2).sum(axis=1))
        sure_in, sure_out = d < a[2][r] - 1e-2, d > a[2][r] + 1e-2
        assert mask[r][sure_in].all() and not mask[r][sure_out].any()


def test_c5_size_population_vs_oracle():
    """BASELINE configs[4] size: 10 000 structures (the shared-memory selection kernel at 160 KB per
    CTA), a small batch of pairs, bit-exact against the numpy oracle; launches repeat bit for bit."""
    g = synth.genome_2mb()
    rad = synth.bead_radii(len(g))
    S = 10000
    rng = np.random.default_rng(5)
    crd = (rng.normal(size=(len(g), S, 3)) * 1500.0).astype(np.float32)
    pi = rng.integers(0, g.n_hap - 1, size=300).astype(np.int32)
    pj = (pi + 1 + rng.integers(0, 40, size=300)).clip(max=g.n_hap - 1).astype(np.int32)
    pw = rng.uniform(0.05, 0.9, size=pi.size)
    pl = np.where(rng.random(pi.size) < 0.5, 0.0, np.round(rng.random(pi.size) * 0.05, 4))
    with Population(S, rad, g.chrom, g.copy_off, g.copy_beads) as pop:
        pop.upload(crd)
        a = pop.actdist(pi, pj, pw, pl, 2.0)
        assert _bits_equal(a, pop.actdist(pi, pj, pw, pl, 2.0)) and a[0].size > 300
    want = ao.actdist_rows(pi, pj, pw, pl, crd, rad, g.chrom, g.copy_index, 2.0)
    assert _bits_equal(a, want)


@pytest.mark.parametrize("shards", [1, 2, 5])
def test_structure_sharded_actdist_is_bit_identical(astep_golden, shards):
    """BASELINE configs[4]: the population sharded by structure, per-pair counts and radix histograms summed over the
    shards (NCCL all-reduce between ranks; here the shards sit on one GPU and their buffers are added) -- the rows must
    equal the reference's own get_actdist output and the single-population kernel, whatever the split"""
    from igm_b200 import parallel
    g = astep_golden
    crd = np.ascontiguousarray(g["coordinates"])
    S = crd.shape[1]
    cuts = np.linspace(0, S, shards + 1).astype(int)
    if shards == 5:
        cuts = np.array([0, 1, 34, 35, 80, S])           # uneven, incl. single-structure shards
    pops = []
    for k in range(shards):
        pop = Population(int(cuts[k + 1] - cuts[k]), g["radii"], g["chrom"], g["copy_off"], g["copy_beads"])
        pop.upload(np.ascontiguousarray(crd[:, cuts[k]:cuts[k + 1], :]))
        pops.append(pop)
    try:
        for chunk in (1 << 17, 97):
            rows = parallel.sharded_actdist(pops, S, g["pair_i"], g["pair_j"], g["pwish"], g["plast"], 2.0, pair_chunk=chunk)
            assert np.array_equal(rows[0], g["row"]) and np.array_equal(rows[1], g["col"])
            assert np.array_equal(rows[2].view(np.uint32), g["dist"].view(np.uint32))
            assert np.array_equal(rows[3].view(np.uint32), g["prob"].view(np.uint32))
    finally:
        for pop in pops:
            pop.close()


def test_structure_sharded_actdist_large_population_vs_full_kernel():
    from igm_b200 import parallel
    g = synth.Genome([9, 7, 6], n_autosomes=2)
    rad = synth.bead_radii(len(g), 1500.0, 0.3)
    S = 2300
    crd = synth.population_territories(g, S, R=1500.0, seed=4)
    iu, ju = np.triu_indices(g.n_hap, 1)
    rng = np.random.default_rng(2)
    pw = rng.uniform(0, 1, iu.size) ** 3
    pl = np.where(rng.random(iu.size) < 0.5, 0.0, rng.uniform(0, 1, iu.size) ** 2)
    with Population(S, rad, g.chrom, g.copy_off, g.copy_beads) as full:
        full.upload(crd)
        want = full.actdist(iu, ju, pw, pl, 2.0)
    cuts = [0, 700, 701, 1500, S]
    pops = []
    for k in range(4):
        pop = Population(cuts[k + 1] - cuts[k], rad, g.chrom, g.copy_off, g.copy_beads)
        pop.upload(np.ascontiguousarray(crd[:, cuts[k]:cuts[k + 1], :]))
        pops.append(pop)
    try:
        got = parallel.sharded_actdist(pops, S, iu, ju, pw, pl, 2.0)
        assert got[0].size > 50
        for a, b in zip(got, want):
            assert np.array_equal(a.view(np.uint32) if a.dtype == np.float32 else a, b.view(np.uint32) if b.dtype == np.float32 else b)
    finally:
        for pop in pops:
            pop.close()
