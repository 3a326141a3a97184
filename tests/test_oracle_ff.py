"""Pins the C restatement of the M-step force field (oracle/ff_oracle.c) against the only
known-answer data the reference ships: the per-structure energies lmp_serial printed for the demo
run (demo/demo_sample_outputs/igm-model.hss.T -> tests/golden/ff_golden.npz)."""
import numpy as np
import pytest

from igm_b200 import synth
from oracle import ff
from helpers import demo_structure, small_system, short_protocol


def _plain_demo(ff_golden, k):
    g = synth.genome_2mb()
    return synth.pack_structure(ff_golden["coordinates"][:, k, :], ff_golden["radii"], g,
                                nucleus_radius=float(ff_golden["nucleus_radius"]))


def test_pair_energy_matches_lammpgen_demo_output(ff_golden):
    """E_pair recomputed from the stored (float32, %g-rounded) coordinates vs what lmp_serial
    reported.  Structures with E ~ 1e-7 are rounding noise and only checked for smallness."""
    for k in range(len(ff_golden["struct_ids"])):
        _, e = ff.single_point(_plain_demo(ff_golden, k))
        want = ff_golden["pair_energies"][k]
        if want > 1.0:
            assert e["pair"] == pytest.approx(want, rel=2e-3), (k, e["pair"], want)
        else:
            assert e["pair"] < 1.0


def test_envelope_energy_matches_lammpgen_demo_output(ff_golden):
    hit = 0
    for k in range(len(ff_golden["struct_ids"])):
        _, e = ff.single_point(_plain_demo(ff_golden, k))
        want = ff_golden["f_envelope0"][k]
        # the dump rounds coordinates to 6 digits (~0.01 at r ~ 5000) while the overshoot t is ~0.1-0.7
        if want > 0.1:
            hit += 1
            assert e["env"][0] == pytest.approx(want, rel=2e-2), (k, e["env"][0], want)
        elif want > 1e-3:
            assert e["env"][0] == pytest.approx(want, rel=0.3), (k, e["env"][0], want)
        else:
            assert e["env"][0] < 1e-2
    assert hit >= 3


def test_total_is_sum_of_terms(ff_golden):
    ps = demo_structure(ff_golden, 0)
    _, e = ff.single_point(ps)
    assert e["total"] == pytest.approx(e["pair"] + e["bond"] + e["env"].sum(), rel=1e-14)
    assert e["bond"] > 0 and e["env"][1] > 0      # the lower bounds and the k<0 envelope are active


def test_forces_are_minus_gradient(ff_golden):
    """central finite differences of the oracle's own energy, on every term at once.  The
    potential is piecewise (bounds switch on/off), so a step of 1/16 can cross a kink of at most
    2*k*h = 0.3 in the force: that is the absolute tolerance.  The k<0 envelope is spherical here:
    for a true ellipsoid the restated force (pull towards the scaled surface point)
    is by construction not the exact gradient of the reported energy."""
    ps = demo_structure(ff_golden, 1, n_extra_bonds=600, damid_axes=(5225.0, 5225.0, 5225.0))
    f, _ = ff.single_point(ps)
    rng = np.random.default_rng(3)
    strongest = np.argsort(-np.abs(f[:3008]).max(axis=1))[:6]
    atoms = list(strongest) + list(rng.choice(3008, 6, replace=False)) + [3009]       # + the centroid
    x0 = ps.xyz.copy()
    checked = 0
    for a in atoms:
        for d in range(3):
            ps.xyz[a, d] = np.float32(x0[a, d] + 0.0625)
            hp = float(ps.xyz[a, d]) - float(x0[a, d])
            ep = ff.single_point(ps)[1]["total"]
            ps.xyz[a, d] = np.float32(x0[a, d] - 0.0625)
            hm = float(x0[a, d]) - float(ps.xyz[a, d])
            em = ff.single_point(ps)[1]["total"]
            ps.xyz[a, d] = x0[a, d]
            num = -(ep - em) / (hp + hm)
            assert num == pytest.approx(f[a, d], rel=1e-3, abs=0.3), (a, d, num, f[a, d])
            checked += abs(f[a, d]) > 30.0
    assert checked >= 6          # enough components where the relative tolerance is the binding one


def test_static_dummy_feels_no_force(ff_golden):
    ps = demo_structure(ff_golden, 2)
    f, _ = ff.single_point(ps)
    assert np.all(f[3008] == 0.0)
    assert np.abs(f[3009]).max() > 0.0            # the centroid moves


def test_bond_prefactor_and_stage_scales(ff_golden):
    ps = demo_structure(ff_golden, 3)
    _, e1 = ff.single_point(ps, 1.0, 1.0, 1.0)
    _, e2 = ff.single_point(ps, 0.5, 1.0, 0.5)
    assert e2["pair"] == pytest.approx(0.5 * e1["pair"], rel=1e-12)
    assert e2["bond"] == pytest.approx(0.5 * e1["bond"], rel=1e-12)
    _, e3 = ff.single_point(ps, 1.0, 0.8, 1.0)
    assert e3["env"][0] > e1["env"][0]            # a smaller envelope violates more


def test_ranpark_known_values():
    """Park-Miller minimal standard: seed 1 -> 16807, 282475249, 1622650073, ... ; the 10000th
    state from seed 1 is 1043618065 (Park & Miller 1988)."""
    v = ff.ranpark(1, 10000)
    m = 2147483647.0
    assert round(v[0] * m) == 16807 and round(v[1] * m) == 282475249 and round(v[2] * m) == 1622650073
    assert round(v[-1] * m) == 1043618065


def test_protocol_lowers_energy_and_satisfies_restraints():
    g, rad, systems = small_system(seed=1)
    ps = systems[0]
    _, e0 = ff.single_point(ps)
    x, res = ff.optimize(ps, short_protocol(md=400, relax=50))
    assert res.status == 0 and np.isfinite(x).all()
    assert res.final_energy < 1e-3 * e0["total"]
    assert res.final_energy == pytest.approx(res.pair_energy + res.bond_energy + res.env_energy[0], rel=1e-12)
    # every bead ends inside the nucleus (+ a small tolerance)
    assert (np.linalg.norm(x[:len(g)], axis=1) + rad <= 1400.0 * 1.02).all()
    assert np.all(x[len(g)] == 0.0)               # the static dummy did not move
