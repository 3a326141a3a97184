"""Force-field parity: the CUDA single point (`run 0`) against the CPU oracle at identical
coordinates.  north_star criterion 1: energies and forces within 1e-5 relative."""
import numpy as np
import pytest

from igm_b200 import synth
from igm_b200.kernel import MStepBatch
from oracle import ff
from helpers import demo_structure, small_system

pytestmark = pytest.mark.gpu

RTOL = 1e-5           # north_star tolerance; the float64 path is expected to sit near 1e-12


def _check(batch, slot, ps, f_ref, e_ref, rtol):
    _, res = batch.fetch(slot)
    f = batch.fetch_forces(slot)
    assert res.final_energy == pytest.approx(e_ref["total"], rel=rtol)
    assert res.pair_energy == pytest.approx(e_ref["pair"], rel=rtol, abs=rtol * abs(e_ref["total"]))
    assert res.bond_energy == pytest.approx(e_ref["bond"], rel=rtol, abs=rtol * abs(e_ref["total"]))
    for j in range(len(e_ref["env"])):
        assert res.env_energy[j] == pytest.approx(e_ref["env"][j], rel=rtol, abs=rtol * abs(e_ref["total"]))
    scale = np.abs(f_ref).max()
    assert np.abs(f - f_ref).max() <= rtol * scale, np.abs(f - f_ref).max() / scale
    return res, f


def test_single_point_f64_all_terms_demo_structures(ff_golden):
    """ten 3008-bead demo structures, every force-field term switched on, in ONE launch"""
    systems = [demo_structure(ff_golden, k) for k in range(len(ff_golden["struct_ids"]))]
    with MStepBatch(0) as b:
        slots = [b.submit(ps) for ps in systems]
        b.single_point()
        for ps, slot in zip(systems, slots):
            f_ref, e_ref = ff.single_point(ps)
            _check(b, slot, ps, f_ref, e_ref, 1e-10)


def test_single_point_matches_lammpgen_demo_energies(ff_golden):
    """the CUDA pair energy against what lmp_serial itself printed for the demo structures"""
    g = synth.genome_2mb()
    with MStepBatch(0) as b:
        slots = [b.submit(synth.pack_structure(ff_golden["coordinates"][:, k, :], ff_golden["radii"], g))
                 for k in range(len(ff_golden["struct_ids"]))]
        b.single_point()
        for k, slot in enumerate(slots):
            _, res = b.fetch(slot)
            want = ff_golden["pair_energies"][k]
            if want > 1.0:
                assert res.pair_energy == pytest.approx(want, rel=2e-3)
            if ff_golden["f_envelope0"][k] > 0.1:
                assert res.env_energy[0] == pytest.approx(ff_golden["f_envelope0"][k], rel=2e-2)


def test_single_point_stage_scales_and_prefactor(ff_golden):
    ps = demo_structure(ff_golden, 4)
    f_ref, e_ref = ff.single_point(ps, 0.5, 1.2, 0.5)
    with MStepBatch(0) as b:
        slot = b.submit(ps)
        b.single_point(0.5, 1.2, 0.5)
        _check(b, slot, ps, f_ref, e_ref, 1e-10)


def test_single_point_f32_md_force_routine(ff_golden):
    """the float32 routine the MD loop uses: same terms, float32 round-off only"""
    systems = [demo_structure(ff_golden, k) for k in (0, 3, 6)]
    with MStepBatch(0) as b:
        slots = [b.submit(ps) for ps in systems]
        b.single_point(f32=True)
        for ps, slot in zip(systems, slots):
            f_ref, e_ref = ff.single_point(ps)
            _check(b, slot, ps, f_ref, e_ref, 2e-4)


def test_no_steric_restraint_means_no_pair_term(ff_golden):
    ps = demo_structure(ff_golden, 0, evfactor=-1.0)
    f_ref, e_ref = ff.single_point(ps)
    assert e_ref["pair"] == 0.0
    with MStepBatch(0) as b:
        slot = b.submit(ps)
        b.single_point()
        res, _ = _check(b, slot, ps, f_ref, e_ref, 1e-10)
        assert res.pair_energy == 0.0


def test_ragged_batch_of_small_systems():
    """structures of different sizes in one launch, incl. a 2-atom one"""
    _, _, a = small_system(seed=2, n_struct=2)
    _, _, b_ = small_system(seed=3, n_struct=1, sizes=(30, 20, 9), n_autosomes=2, n_hic=10)
    g1 = synth.Genome([1], n_autosomes=0)
    tiny = synth.pack_structure(np.array([[10.0, 0, 0]], dtype=np.float32), np.array([50.0], dtype=np.float32), g1,
                                nucleus_radius=55.0)
    systems = a + b_ + [tiny]
    with MStepBatch(0) as b:
        slots = [b.submit(ps) for ps in systems]
        b.single_point()
        for ps, slot in zip(systems, slots):
            f_ref, e_ref = ff.single_point(ps)
            _, res = b.fetch(slot)
            f = b.fetch_forces(slot)
            assert res.final_energy == pytest.approx(e_ref["total"], rel=1e-10, abs=1e-12)
            assert np.abs(f - f_ref).max() <= 1e-10 * max(np.abs(f_ref).max(), 1e-30)


def test_dense_start_widens_the_neighbour_table():
    """territory starts pack > 96 beads inside the list cutoff: the run must widen, not fail"""
    g = synth.Genome([400], n_autosomes=0)
    rad = synth.bead_radii(len(g), 1500.0, 0.4)
    xyz = (np.random.default_rng(1).normal(size=(400, 3)) * 120.0).astype(np.float32)
    ps = synth.pack_structure(xyz, rad, g, nucleus_radius=1500.0)
    f_ref, e_ref = ff.single_point(ps)
    with MStepBatch(0) as b:
        slot = b.submit(ps)
        b.single_point()
        _check(b, slot, ps, f_ref, e_ref, 1e-10)


def test_errors_are_reported_not_thrown():
    with MStepBatch(0) as b:
        with pytest.raises(RuntimeError, match="nothing submitted"):
            b.single_point()
        g, rad, s = small_system()
        ps = s[0]
        ps.bond_i[0] = 10 ** 6
        with pytest.raises(RuntimeError, match="out of range"):
            b.submit(ps)
