"""FISH A-step: restatement and CUDA kernel against the reference's own helper outputs."""
import os

import numpy as np
import pytest

from conftest import GOLDEN
from oracle import fish_oracle as fo


@pytest.fixture(scope="module")
def fg():
    return np.load(os.path.join(GOLDEN, "fish_golden.npz"))


def _copy_index(g):
    off, beads = g["copy_off"], g["copy_beads"]
    return {h: [int(b) for b in beads[off[h]:off[h + 1]]] for h in range(len(off) - 1)}


def _check(got, fg, prefix):
    lo, hi, ra, rb = got
    assert np.array_equal(lo.astype(np.float64), fg[prefix + "min"])
    assert np.array_equal(hi.astype(np.float64), fg[prefix + "max"])
    assert np.array_equal(ra, fg[prefix + "imin"]) and np.array_equal(rb, fg[prefix + "imax"])


def test_oracle_matches_reference(astep_golden, fg):
    g = astep_golden
    ci = _copy_index(g)
    _check(fo.fish_ranks(g["coordinates"], ci, fg["pair_i"], fg["pair_j"]), fg, "p")
    _check(fo.fish_ranks(g["coordinates"], ci, fg["probes"]), fg, "r")


@pytest.mark.gpu
def test_gpu_matches_reference_and_oracle(astep_golden, fg):
    from igm_b200.astep import Population
    g = astep_golden
    with Population(g["coordinates"].shape[1], g["radii"], g["chrom"], g["copy_off"], g["copy_beads"]) as pop:
        pop.upload(np.ascontiguousarray(g["coordinates"]))
        _check(pop.fish_ranks(fg["pair_i"], fg["pair_j"]), fg, "p")
        _check(pop.fish_ranks(fg["probes"]), fg, "r")
        # mixed batch, and a haploid (single-copy) locus is refused like the reference's IndexError
        hi = np.concatenate([fg["pair_i"][:3], fg["probes"][:2]])
        hj = np.concatenate([fg["pair_j"][:3], [-1, -1]])
        lo, _, ra, _ = pop.fish_ranks(hi, hj)
        assert np.array_equal(ra[:3], fg["pimin"][:3]) and np.array_equal(ra[3:], fg["rimin"][:2])
        single = int(np.nonzero(np.diff(g["copy_off"]) == 1)[0][0])
        with pytest.raises(RuntimeError, match="fewer than two copies"):
            pop.fish_ranks([single])


@pytest.mark.gpu
def test_gpu_large_population_vs_oracle():
    from igm_b200 import synth
    from igm_b200.astep import Population
    g = synth.Genome([9, 7, 6], n_autosomes=3)
    rad = synth.bead_radii(len(g), 1500.0, 0.3)
    crd = synth.population_territories(g, 1234, R=1500.0, seed=8)
    hi = np.array([0, 5, 12, 20], dtype=np.int32)
    hj = np.array([3, -1, 21, -1], dtype=np.int32)
    want = fo.fish_ranks(crd, g.copy_index, hi, hj)
    with Population(1234, rad, g.chrom, g.copy_off, g.copy_beads) as pop:
        pop.upload(crd)
        got = pop.fish_ranks(hi, hj)
    for a, w in zip(got, want):
        assert np.array_equal(a, w)
