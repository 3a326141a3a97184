"""DamID (lamina) A-step: the numpy restatement and the CUDA kernels against outputs of the
reference's own get_damid_actdist and Damid._apply_envelope (tests/golden/damid_golden.npz)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN
from oracle import damid_oracle as do


@pytest.fixture(scope="module")
def dg():
    return np.load(os.path.join(GOLDEN, "damid_golden.npz"))


def _copy_index(g):
    off, beads = g["copy_off"], g["copy_beads"]
    return {h: [int(b) for b in beads[off[h]:off[h + 1]]] for h in range(len(off) - 1)}


SHAPES = ["sphere", "ellipsoid"]


def _semi(dg, shape):
    return (float(dg["sphere_param"]),) * 3 if shape == "sphere" else tuple(dg["ellipsoid_param"])


@pytest.mark.parametrize("shape", SHAPES)
def test_oracle_rows_and_members_match_reference(astep_golden, dg, shape):
    g = astep_golden
    param = float(dg["sphere_param"]) if shape == "sphere" else list(dg["ellipsoid_param"])
    cr = float(dg[shape + "_cr"])
    bead, dist, prob = do.damid_rows(dg["loc"], dg["pwish"], dg["plast"], g["coordinates"], g["radii"], _copy_index(g), cr, shape, param)
    assert np.array_equal(bead, dg[shape + "_bead"])
    assert np.array_equal(dist.view(np.uint32), dg[shape + "_dist"].view(np.uint32))
    assert np.array_equal(prob.view(np.uint32), dg[shape + "_prob"].view(np.uint32))
    mask = do.damid_assign(g["coordinates"], g["radii"], bead, dist, cr, _semi(dg, shape))
    for s in g["assign_structs"]:
        assert sorted(bead[mask[:, s]]) == list(dg["%s_members_%d" % (shape, s)])


@pytest.mark.gpu
@pytest.mark.parametrize("shape", SHAPES)
def test_gpu_damid_bit_exact(astep_golden, dg, shape):
    from igm_b200.astep import Population, unpack_bits
    g = astep_golden
    param = float(dg["sphere_param"]) if shape == "sphere" else list(dg["ellipsoid_param"])
    cr = float(dg[shape + "_cr"])
    with Population(g["coordinates"].shape[1], g["radii"], g["chrom"], g["copy_off"], g["copy_beads"]) as pop:
        pop.upload(np.ascontiguousarray(g["coordinates"]))
        bead, dist, prob = pop.damid_actdist(dg["loc"], dg["pwish"], dg["plast"], cr, shape, param)
        assert np.array_equal(bead, dg[shape + "_bead"])
        assert np.array_equal(dist.view(np.uint32), dg[shape + "_dist"].view(np.uint32))
        assert np.array_equal(prob.view(np.uint32), dg[shape + "_prob"].view(np.uint32))
        bits = pop.damid_assign(bead, dist, cr, _semi(dg, shape))
    mask = unpack_bits(bits, g["coordinates"].shape[1])
    assert np.array_equal(mask, do.damid_assign(g["coordinates"], g["radii"], bead, dist, cr, _semi(dg, shape)))
    for s in g["assign_structs"]:
        assert sorted(bead[mask[:, s]]) == list(dg["%s_members_%d" % (shape, s)])


@pytest.mark.gpu
@pytest.mark.parametrize("n_struct", [1, 77, 1000])
def test_gpu_damid_vs_oracle_synthetic(n_struct):
    from igm_b200 import synth
    from igm_b200.astep import Population
    g = synth.Genome([9, 7, 6, 4, 3], n_autosomes=3)
    rad = synth.bead_radii(len(g), 1500.0, 0.3)
    rad[::5] *= np.float32(0.9)                          # several distinct radii
    crd = synth.population_territories(g, n_struct, R=1500.0, seed=n_struct)
    rng = np.random.default_rng(n_struct)
    loc = np.arange(g.n_hap, dtype=np.int32)
    pw = np.round(rng.random(loc.size), 3)
    pl = np.where(rng.random(loc.size) < 0.5, 0.0, np.round(rng.random(loc.size) * pw, 5))
    for shape, param, cr in (("sphere", 1500.0, 0.05), ("ellipsoid", [1600.0, 1450.0, 1300.0], 0.1)):
        want = do.damid_rows(loc, pw, pl, crd, rad, g.copy_index, cr, shape, param)
        with Population(n_struct, rad, g.chrom, g.copy_off, g.copy_beads) as pop:
            pop.upload(crd)
            got = pop.damid_actdist(loc, pw, pl, cr, shape, param)
            semi = (param,) * 3 if shape == "sphere" else tuple(param)
            bits = pop.damid_assign(got[0], got[1], cr, semi)
        for a, w in zip(got, want):
            assert np.array_equal(a.view(np.uint32), w.view(np.uint32))
        from oracle import astep_oracle as ao
        assert np.array_equal(bits, ao.pack_bits(do.damid_assign(crd, rad, got[0], got[1], cr, semi)))
