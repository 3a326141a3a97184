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


def test_sprite_centroids_and_bonds_equal_the_reference(astep_golden, rg):
    """Sprite._apply (igm/restraints/sprite.py:41-69) for a block of structures against the reference's own Model"""
    from igm_b200 import sprite
    g = astep_golden
    crd = np.ascontiguousarray(g["coordinates"])
    n_bead = crd.shape[0]
    indptr, sel = rg["sprite_indptr"], rg["sprite_selected"]
    beads = [sel[indptr[c]:indptr[c + 1]] for c in range(len(indptr) - 1)]
    atoms, bonds = sprite.centroid_restraints(crd, g["radii"], rg["sprite_assignment"], beads, 0, n_bead,
                                              float(rg["sprite_volume_occupancy"]), float(rg["sprite_k"]))
    aoff, xyz, arad, akind = atoms
    boff, bi, bj, bk, br0, bs = bonds
    for s in rg["structs"]:
        s = int(s)
        want_f, want_c = rg["sprite_forces_%d" % s], rg["sprite_centroids_%d" % s]
        assert np.array_equal(xyz[aoff[s]:aoff[s + 1]].view(np.uint32), want_c.view(np.uint32))
        sl = slice(boff[s], boff[s + 1])
        got = np.stack([bi[sl], bj[sl], br0[sl], bk[sl]], axis=1).astype(np.float64)
        assert np.array_equal(got, want_f)
        assert (akind[aoff[s]:aoff[s + 1]] == 2).all() and (arad[aoff[s]:aoff[s + 1]] == 0).all() and (bs[sl] == 0).all()
    assert aoff[-1] == np.isin(rg["sprite_assignment"], np.arange(crd.shape[1])).sum()


def test_gibbs_assignment_replays_the_reference_random_stream():
    """SpriteAssignmentStep.reduce (:176-272) draws from the global numpy state; assign_clusters takes a RandomState and
    must consume it with the same calls in the same order: np.random.seed(k) + the restatement == RandomState(k) + product"""
    from igm_b200 import sprite
    from oracle import sprite_oracle as so
    rng = np.random.default_rng(3)
    n_cl, keep, S = 137, 12, 40
    rg2 = np.sort(rng.uniform(100.0, 5e5, (n_cl, keep)).astype(np.float32), axis=1)
    idx = np.stack([rng.choice(S, keep, replace=False) for _ in range(n_cl)]).astype(np.int32)
    rg2[[5, 77]] = -1.0; idx[[5, 77]] = -1                                         # clusters skipped by the A-step
    for seed, bs, kT in ((1, 10, 100.0), (9, 7, 30.0)):
        want = so.reduce_assignment(rg2, idx, S, bs, kT, seed)
        got = sprite.assign_clusters(rg2, idx, S, np.random.RandomState(seed), batch_size=bs, kT=kT)
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
        assert got[0][5] == -1 and np.bincount(got[0][got[0] >= 0], minlength=S).max() <= 12
