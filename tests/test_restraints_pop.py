"""Per-structure restraints (FISH bonds, DamID envelope members): the reference's own force lists (golden), the
per-structure restatement (oracle/restraints_oracle.py) and the population-wide builder of the product
(igm_b200/restraints_pop.py) must agree item for item.  Host logic: runs without a GPU."""
import os

import numpy as np
import pytest

from conftest import GOLDEN
from igm_b200 import restraints_pop as rp
from oracle import restraints_oracle as ro


@pytest.fixture(scope="module")
def rg():
    return np.load(os.path.join(GOLDEN, "restraints_golden.npz"))


def _copy_index(g):
    off, beads = g["copy_off"], g["copy_beads"]
    return {h: [int(b) for b in beads[off[h]:off[h + 1]]] for h in range(len(off) - 1)}


def test_oracle_reproduces_the_reference_force_lists(astep_golden, rg):
    g = astep_golden
    ci = _copy_index(g)
    dg = np.load(os.path.join(GOLDEN, "damid_golden.npz"))
    for s in rg["structs"]:
        s = int(s)
        dummies = rg["dummies_%d" % s]
        assert len(dummies) == 2                     # Damid's centre, then Fish's (damid.py:120, fish.py:119)
        got = ro.fish_forces(g["coordinates"][:, s, :], ci, int(dummies[1]), "rRpP", float(rg["tol"]), float(rg["kspring"]), s,
                             probes=rg["probes"], radial_min=rg["radial_min"], radial_max=rg["radial_max"],
                             pairs=rg["pairs"], pair_min=rg["pair_min"], pair_max=rg["pair_max"])
        want = rg["fish_forces_%d" % s]
        assert len(got) == want.shape[0]
        assert np.array_equal(np.array(got, dtype=np.float64), want)
        mem = ro.damid_members(g["coordinates"][:, s, :], g["radii"], dg["ellipsoid_bead"], dg["ellipsoid_dist"],
                               float(dg["ellipsoid_cr"]), dg["ellipsoid_param"])
        assert mem == [int(v) for v in rg["damid_members_%d" % s]]


def test_population_builder_equals_the_reference_for_every_structure(astep_golden, rg):
    g = astep_golden
    crd = np.ascontiguousarray(g["coordinates"])
    S = crd.shape[1]
    center = int(rg["dummies_0"][1])
    off, bi, bj, bk, br0, bs = rp.fish_bonds(crd, g["copy_off"], g["copy_beads"], center, "rRpP", float(rg["tol"]), float(rg["kspring"]),
                                            probes=rg["probes"], radial_min=rg["radial_min"], radial_max=rg["radial_max"],
                                            pairs=rg["pairs"], pair_min=rg["pair_min"], pair_max=rg["pair_max"])
    assert off.size == S + 1 and off[-1] == bi.size
    for s in rg["structs"]:                          # against the reference itself
        s = int(s)
        want = rg["fish_forces_%d" % s]
        sl = slice(off[s], off[s + 1])
        got = np.stack([bs[sl], bi[sl], bj[sl], br0[sl], bk[sl]], axis=1).astype(np.float64)
        assert np.array_equal(got, want)
    ci = _copy_index(g)
    for s in (1, 42, 77):                            # and against the restatement on structures the golden file does not hold
        want = np.array(ro.fish_forces(crd[:, s, :], ci, center, "rRpP", float(rg["tol"]), float(rg["kspring"]), s, probes=rg["probes"],
                                       radial_min=rg["radial_min"], radial_max=rg["radial_max"], pairs=rg["pairs"],
                                       pair_min=rg["pair_min"], pair_max=rg["pair_max"]), dtype=np.float64)
        sl = slice(off[s], off[s + 1])
        assert np.array_equal(np.stack([bs[sl], bi[sl], bj[sl], br0[sl], bk[sl]], axis=1).astype(np.float64), want)


def test_fish_targets_gather(rg):
    fg = np.load(os.path.join(GOLDEN, "fish_golden.npz"))
    assert np.array_equal(rp.fish_targets(fg["rimin"][:12], rg["radial_min_sorted"]), rg["radial_min"])
    assert np.array_equal(rp.fish_targets(fg["pimax"][:15], rg["pair_max_sorted"]), rg["pair_max"])
    with pytest.raises(ValueError):
        rp.fish_targets(fg["rimin"][:12], rg["radial_min_sorted"][:, :50])


def test_subset_of_restraint_types_and_empty(astep_golden, rg):
    g = astep_golden
    crd = np.ascontiguousarray(g["coordinates"][:, :7, :])
    off, bi, *_ = rp.fish_bonds(crd, g["copy_off"], g["copy_beads"], 702, "", 10.0, 1.0)
    assert off.tolist() == [0] * 8 and bi.size == 0
    off, bi, bj, bk, br0, bs = rp.fish_bonds(crd, g["copy_off"], g["copy_beads"], 702, "R", 10.0, 1.0, probes=rg["probes"],
                                            radial_max=rg["radial_max"][:, :7])
    assert off[1] == 3 * len(rg["probes"]) and set(bi.tolist()) == {702}
    env = rp.damid_envelope([1, 2, 3], np.zeros((3, 1), dtype=np.uint32), (5000.0, 4000.0, 3000.0), 0.05, 0.3)
    assert env[0] == (4750.0, 3800.0, 2850.0, -0.3)
