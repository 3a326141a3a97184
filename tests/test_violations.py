"""M10 violation ratios: restatement and CUDA kernels against the reference's Force classes
(tests/golden/viol_golden.npz from HarmonicUpperBound/LowerBound.getViolationRatio and
EllipticEnvelope.getViolationRatios)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN
from igm_b200 import _abi
from igm_b200.kernel import MStepBatch, envelope_ratios, violation_histogram
from oracle import viol_oracle as vo


@pytest.fixture(scope="module")
def vg():
    return np.load(os.path.join(GOLDEN, "viol_golden.npz"))


def _xyz(astep_golden, vg):
    return astep_golden["coordinates"][:int(vg["n_bead"]), int(vg["struct"])], astep_golden["radii"][:int(vg["n_bead"])]


def test_oracle_matches_reference(astep_golden, vg):
    xyz, rad = _xyz(astep_golden, vg)
    r = vo.bond_ratios(xyz, vg["bond_i"], vg["bond_j"], vg["bond_r0"], vg["bond_k"], vg["bond_style"])
    assert np.array_equal(r, vg["bond_ratio"]) and np.count_nonzero(r) > 100
    for e in (0, 1):
        u = vo.envelope_u(xyz, rad, vg["env%d_members" % e], vg["env%d_semi" % e], float(vg["env%d_k" % e]))
        assert np.array_equal(envelope_ratios(u, float(vg["env%d_k" % e]), float(vg["env%d_scale" % e])), vg["env%d_ratio" % e])


def test_histogram_semantics():
    H, edges = violation_histogram([0.0, 0.004, 0.5, 1.0, 1.5, 7.0])
    assert H.sum() == 6 and H[-1] == 2 and H[0] == 2 and H[50] == 1 and H[99] == 1 and len(edges) == 102


@pytest.mark.gpu
def test_gpu_ratios_bit_exact(astep_golden, vg):
    xyz, rad = _xyz(astep_golden, vg)
    n = len(rad)
    members = [vg["env0_members"], vg["env1_members"]]
    ps = _abi.PackedSystem(xyz, rad, np.zeros(n, dtype=np.uint8), vg["bond_i"], vg["bond_j"], vg["bond_k"], vg["bond_r0"],
                           vg["bond_style"], np.array([list(vg["env0_semi"]) + [float(vg["env0_k"])],
                                                       list(vg["env1_semi"]) + [float(vg["env1_k"])]]),
                           [0, len(members[0]), len(members[0]) + len(members[1])], np.concatenate(members), 1.0)
    with MStepBatch(0) as b:
        slot = b.submit(ps)
        br, eu = b.violation_ratios(slot)
    assert np.array_equal(br, vg["bond_ratio"])
    n0 = len(members[0])
    assert np.array_equal(envelope_ratios(eu[:n0], float(vg["env0_k"]), float(vg["env0_scale"])), vg["env0_ratio"])
    assert np.array_equal(eu[n0:], vo.envelope_u(xyz, rad, members[1], vg["env1_semi"], float(vg["env1_k"])))
    assert np.array_equal(envelope_ratios(eu[n0:], float(vg["env1_k"]), float(vg["env1_scale"])), vg["env1_ratio"])


@pytest.mark.gpu
def test_gpu_ratios_follow_the_optimised_coordinates():
    from helpers import small_system, short_protocol
    g, rad, systems = small_system(seed=31, n_struct=3)
    with MStepBatch(0) as b:
        slots = [b.submit(ps) for ps in systems]
        before = [b.violation_ratios(s)[0] for s in slots]
        b.run(short_protocol(md=300, relax=50))
        for ps, slot, old in zip(systems, slots, before):
            xyz, _ = b.fetch(slot)
            br, eu = b.violation_ratios(slot)
            assert np.array_equal(br, vo.bond_ratios(xyz, ps.bond_i, ps.bond_j, ps.bond_r0, ps.bond_k, ps.bond_style))
            assert np.count_nonzero(br > 0.05) <= np.count_nonzero(old > 0.05)
            assert np.array_equal(eu, vo.envelope_u(xyz, ps.radius, ps.env_members, ps.env_abck[0][:3], ps.env_abck[0][3]))
