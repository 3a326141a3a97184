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
                                                      (1500, 5, False), (2100, 6, False), (1000, 7, True)])
def test_actdist_and_assign_vs_oracle_synthetic(n_struct, seed, force_smem, monkeypatch):
    """ragged structure counts (not multiples of 32 / 256), X/Y single-copy bins, plast in {0, <1, 1};
    covers the register-resident kernel (<= 2048 structures) and the shared-memory one"""
    if force_smem:
        monkeypatch.setenv("IGMK_ACTDIST_SMEM", "1")
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
