"""The Model -> flat-array packer and the protocol builder against what the reference's own LAMMPS
front end wrote for the same model (tests/golden/lammps_golden.json, made by make_golden.py with
LammpsModel / create_lammps_data / create_lammps_script)."""
import json
import math
import os

import numpy as np
import pytest

from conftest import GOLDEN
from igm_b200 import _abi
from igm_b200.pack import pack_model, script_seed
from oracle import ff


class _P(object):
    def __init__(self, pos, r, ptype):
        self.pos = np.array(pos, dtype=np.float32); self.r = np.float32(r); self.ptype = ptype


class _F(object):
    def __init__(self, d):
        self.__dict__.update(d)
        if "semiaxes" in d:
            self.semiaxes = np.array(d["semiaxes"])


class _Model(object):
    pass


@pytest.fixture(scope="module")
def golden():
    return json.load(open(os.path.join(GOLDEN, "lammps_golden.json")))


@pytest.fixture(scope="module")
def model(golden):
    m = _Model()
    m.id = golden["model_id"]
    pp = golden["particles"]
    m.particles = [_P(p, r, t) for p, r, t in zip(pp["pos"], pp["r"], pp["ptype"])]
    m.forces = [_F(f) for f in golden["forces"]]
    return m


def test_seed_formula(golden):
    # velocity lines carry seed + stage (lammps.py:322-323,335)
    seed = script_seed({"seed": golden["seed"]}, golden["model_id"], golden["step_no"])
    assert golden["script_velocity_lines"][0].split()[-1] == str(seed)
    assert golden["script_velocity_lines"][2].split()[-1] == str(seed + 1)


def test_atoms_and_bonds_match_the_data_file(golden, model):
    ps = pack_model(model, seed=1)
    imap = golden["imap"]
    atoms = {a[0]: a for a in golden["atoms"]}                 # id -> (id, mol, type, x, y, z), 1-based ids
    for i, p in enumerate(model.particles):
        a = atoms[imap[i] + 1]
        assert np.allclose(ps.xyz[i], a[3:6], rtol=0, atol=0)
    # bonds: same multiset of (style, k, r0, endpoints) with dummy clones folded back onto the dummy
    coeff = {c[0]: c for c in golden["bond_coeffs"]}
    back = {}
    for i, aid in enumerate(imap):
        back[aid + 1] = i
    # all static dummies sit at the origin and are interchangeable: LAMMPS even clones them
    # (<= 20 bonds per atom, lammps_model.py:186,303-312), so they are compared as one node, -1
    is_dummy = [p.ptype == 1 for p in model.particles]
    canon = lambda i: -1 if (i is None or is_dummy[i]) else int(i)
    want = []
    for _, btype, a, b in golden["bonds"]:
        st, k, r0 = coeff[btype][1:]
        ia, ib = canon(back.get(a)), canon(back.get(b))         # clones of a static dummy have no particle
        want.append((st, round(k, 9), round(r0, 6), min(ia, ib), max(ia, ib)))
    got = []
    for i, j, k, r0, st in zip(ps.bond_i, ps.bond_j, ps.bond_k, ps.bond_r0, ps.bond_style):
        got.append(("harmonic_upper_bound" if st == _abi.BOND_UPPER else "harmonic_lower_bound",
                    round(float(k), 9), round(float(r0), 6), min(canon(i), canon(j)), max(canon(i), canon(j))))
    assert sorted(got) == sorted(want)
    assert ps.evfactor == 1.0


def test_envelopes_match_the_script(golden, model):
    ps = pack_model(model)
    # stage 0 scales the semi-axes by envelope_factors[0] = 1.2 (lammps.py:298-299)
    line0 = golden["script_envelope_lines"][0].split()
    line1 = golden["script_envelope_lines"][1].split()
    assert [float(v) for v in line0[4:8]] == pytest.approx(list(ps.env_abck[0][:3] * 1.2) + [ps.env_abck[0][3]])
    assert [float(v) for v in line1[4:8]] == pytest.approx(list(ps.env_abck[1][:3] * 1.2) + [ps.env_abck[1][3]])
    assert ps.env_abck[1][3] < 0                                # the DamID-style envelope
    assert list(ps.env_members[ps.env_off[1]:ps.env_off[2]]) == [5, 6, 7]


def test_pair_coefficients_are_what_the_oracle_uses(golden, model):
    """PairIJ Coeffs: A = ((ri+rj)/pi)^2 * evfactor, rc = ri+rj for bead-bead, 0 0 otherwise (lammps.py:117-137)"""
    ps = pack_model(model)
    types = {}
    for a in golden["atoms"]:
        types[a[0]] = a[2]
    radius = {u[0]: u[2] for u in golden["user"]}
    rad_of_type = {}
    for aid, t in types.items():
        rad_of_type.setdefault(t, radius[aid])
    for ti, tj, A, rc in golden["pair_coeffs"]:
        ri, rj = rad_of_type[ti], rad_of_type[tj]
        if A > 0:
            assert rc == pytest.approx(ri + rj, rel=1e-6)
            assert A == pytest.approx(((ri + rj) / math.pi) ** 2, rel=1e-5)
    # and the oracle's single point on two overlapping beads reproduces A (1 + cos(pi r / rc))
    two = _abi.PackedSystem([[0, 0, 0], [100.0, 0, 0]], [80.0, 70.0], [0, 0], [], [], [], [], [],
                            np.zeros((0, 4)), [0], [], 1.0)
    _, e = ff.single_point(two)
    A = (150.0 / math.pi) ** 2
    assert e["pair"] == pytest.approx(A * (1 + math.cos(math.pi * 100.0 / 150.0)), rel=1e-12)


def test_protocol_matches_the_script(golden):
    prot = _abi.make_protocol(golden["optimizer_options"])
    runs = [int(l.split()[1]) for l in golden["script_run_lines"] if l.startswith("run")]
    assert prot.n_stage == 4 and prot.relax_mdsteps == runs[0]
    assert [prot.mdsteps[k] for k in range(4)] == runs[1::2]
    resc = golden["script_rescale_lines"]                       # relax, main, relax, main ...
    for k in range(4):
        main = resc[2 * k + 1].split()
        assert float(main[5]) == prot.tstart[k] and float(main[6]) == prot.tstop[k]
    mini = [l for l in golden["script_run_lines"] if l.startswith("minimize")][0].split()
    assert (float(mini[1]), float(mini[2]), int(mini[3]), int(mini[4])) == (prot.etol, prot.ftol, prot.max_cg_iter, prot.max_cg_eval)
    ts = [l for l in golden["script_misc_lines"] if l.startswith("timestep")][0]
    assert float(ts.split()[1]) == prot.timestep
    evf = [float(l.split()[-1]) for l in golden["script_misc_lines"] if l.startswith("variable evprefactor")]
    assert evf == [prot.evfactor[k] for k in range(4)]


def test_unknown_particle_type_and_voxel_envelope_are_rejected(model):
    m = _Model(); m.id = 1
    m.particles = [_P([0, 0, 0], 1.0, 7)]; m.forces = []
    with pytest.raises(ValueError):
        pack_model(m)
    m.particles = [_P([0, 0, 0], 1.0, 0)]
    m.forces = [_F({"ftype": 4, "k": 1.0})]
    with pytest.raises(NotImplementedError):
        pack_model(m)
