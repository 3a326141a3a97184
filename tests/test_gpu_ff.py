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


MD_FORCE_RTOL = 2e-4   # float32 sums of ~30 terms + MUFU rsqrt / sin / rcp approximations (~1e-6 each), relative to max |F|


@pytest.mark.parametrize("stage,general", [(0, False), (3, False), (1, True)])
def test_md_kernel_force_routine_matches_oracle(ff_golden, stage, general, monkeypatch):
    """The force loop k_md2 itself runs (oriented bond lists, inboxes, neighbour tables, float32, hardware
    approximations) against the double-precision oracle: 3008 beads, every term on, the stage's pair / envelope
    scales.  Evaluated through the full bond lists and again through the candidate lists.  The demo beads all have one
    radius, which selects the pair loop specialised for that; `general` forces the loop for mixed radii."""
    from helpers import short_protocol
    if general:
        monkeypatch.setenv("IGMK_GENERAL_PAIRS", "1")
    prot = short_protocol()
    systems = [demo_structure(ff_golden, k) for k in (0, 3, 6, 8)]
    ev, env = prot.evfactor[stage], prot.envfactor[stage]
    with MStepBatch(0) as b:
        slots = [b.submit(ps) for ps in systems]
        b.md_forces(prot, stage)
        for ps, slot in zip(systems, slots):
            f_ref, _ = ff.single_point(ps, ev, env, 1.0)
            f = b.fetch_forces(slot)
            scale = np.abs(f_ref).max()
            assert np.abs(f - f_ref).max() <= MD_FORCE_RTOL * scale, np.abs(f - f_ref).max() / scale
            # per atom: relative to that atom's own force where it is not tiny
            big = np.linalg.norm(f_ref, axis=1) > 1e-3 * scale
            rel = np.linalg.norm(f - f_ref, axis=1)[big] / np.linalg.norm(f_ref, axis=1)[big]
            assert rel.max() <= 2e-3, rel.max()


def test_md_kernel_force_routine_small_and_hub_systems():
    """more bonds on one atom than an inbox holds (SYM items at both ends), lower bounds as the most frequent type,
    a structure without bonds"""
    from helpers import short_protocol
    prot = short_protocol()
    g, rad, sys_ = small_system(seed=4, n_struct=2, n_hic=300)
    n = len(g)
    rng = np.random.default_rng(9)
    ps0 = sys_[0]
    # hub: atom 3 bonded to everyone, twice; mostly lower bounds of one type
    hi = np.concatenate([np.full(2 * (n - 1), 3), rng.integers(0, n, 200)]).astype(np.int32)
    hj = np.concatenate([np.delete(np.arange(n), 3)] * 2 + [rng.integers(0, n, 200)]).astype(np.int32)
    keep = hi != hj
    hi, hj = hi[keep], hj[keep]
    from igm_b200._abi import PackedSystem, BOND_LOWER
    hub = PackedSystem(ps0.xyz, ps0.radius, ps0.kind, hi, hj, np.full(hi.size, 0.7), np.full(hi.size, 3.0 * float(rad[0])),
                       np.full(hi.size, BOND_LOWER, dtype=np.uint8), ps0.env_abck, ps0.env_off, ps0.env_members, 1.0, seed=3)
    bare = synth.pack_structure(sys_[1].xyz[:n], rad, g, nucleus_radius=1400.0)       # polymer bonds only
    systems = [sys_[0], sys_[1], hub, bare]
    with MStepBatch(0) as b:
        slots = [b.submit(ps) for ps in systems]
        b.md_forces(prot, 1)
        for ps, slot in zip(systems, slots):
            f_ref, _ = ff.single_point(ps, prot.evfactor[1], prot.envfactor[1], 1.0)
            f = b.fetch_forces(slot)
            scale = np.abs(f_ref).max()
            assert np.abs(f - f_ref).max() <= MD_FORCE_RTOL * scale, np.abs(f - f_ref).max() / scale


def test_single_point_f32_md_force_routine(ff_golden):
    """the team-per-atom routine of mstep.cuh in float32: same terms, float32 round-off only"""
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


def test_odd_atom_count():
    """an odd number of atoms puts the 16-bit cell cursors of the float64 kernels next to the last atom slot"""
    for sizes in ((13, 11, 8, 6), (14, 11, 8, 6)):
        g = synth.Genome(list(sizes), n_autosomes=3)
        rad = synth.bead_radii(len(g), 1400.0, 0.3)
        rng = np.random.default_rng(7)
        ps = synth.pack_structure(synth.territories(g, 1400.0, rng), rad, g, nucleus_radius=1400.0)
        f_ref, e_ref = ff.single_point(ps)
        with MStepBatch(0) as b:
            slot = b.submit(ps)
            b.single_point()
            _check(b, slot, ps, f_ref, e_ref, 1e-10)


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
