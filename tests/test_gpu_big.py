"""Structures beyond one SM's shared memory (BASELINE config 4 territory: 200 kb genomes): the cluster-per-structure
MD kernel (md_cluster.cuh: atoms split over the CTAs of a thread-block cluster, partner positions through distributed
shared memory), the one-CTA global-memory kernel it replaced (md_big.cuh, kept behind IGMK_FORCE_BIG) and the
global-state mode of the float64 single point / CG."""
import numpy as np
import pytest

from igm_b200 import synth
from igm_b200.kernel import MStepBatch
from oracle import ff
from helpers import small_system, short_protocol

pytestmark = pytest.mark.gpu


def _big_system(n_chain=10, chain=950, seed=3):
    g = synth.Genome([chain] * n_chain, n_autosomes=0)          # 9500 beads: K = 10 slots, above every 1-CTA limit
    R = 6000.0
    rad = synth.bead_radii(len(g), R, 0.3)
    rng = np.random.default_rng(seed)
    xyz = synth.territories(g, R, rng)
    hi = rng.integers(0, len(g), 3000)
    hj = (hi + rng.integers(1, len(g) - 1, 3000)) % len(g)
    return g, rad, synth.pack_structure(xyz, rad, g, hi, hj, np.full(3000, 4.0 * float(rad[0])), nucleus_radius=R, seed=5)


def test_big_md_kernel_follows_the_small_one(monkeypatch):
    """the same small structures through k_md2 and, forced, through k_md_big: the same trajectory up to
    float32 round-off (different instruction selection: 1 ulp after 2 steps, then chaos takes over),
    and the big kernel reproduces itself bit for bit"""
    g, rad, systems = small_system(seed=41, n_struct=3)
    prot = short_protocol(md=5, relax=0)
    prot.n_stage = 1
    prot.skip_minimize = 1
    out = []
    for force in (False, True, True):
        if force:
            monkeypatch.setenv("IGMK_FORCE_BIG", "1")
        with MStepBatch(0) as b:
            slots = [b.submit(ps) for ps in systems]
            b.run(prot)
            out.append([b.fetch(s) for s in slots])
    for (a, ra), (c, rc), (e, _) in zip(*out):
        assert np.abs(a - c).max() <= 1e-3 and ra.n_rebuilds == rc.n_rebuilds
        assert np.array_equal(c.view(np.uint32), e.view(np.uint32))


@pytest.mark.parametrize("cl", [2, 4, 8])
def test_cluster_md_kernel_follows_the_single_cta_one(monkeypatch, cl):
    """the same structures through k_md2 (one CTA) and, forced, through k_md2c on clusters of 2 / 4 / 8 CTAs: the same
    trajectory up to float32 round-off (sums over bonds and CTAs are taken in another order), the same rebuild count,
    and the cluster kernel reproduces itself bit for bit"""
    g, rad, systems = small_system(seed=41, n_struct=3)
    from helpers import demo_structure
    ffg = np.load(__import__("os").path.join(__import__("conftest").GOLDEN, "ff_golden.npz"))
    systems = systems + [demo_structure(ffg, 2)]              # 3010 atoms with every kind of term
    prot = short_protocol(md=6, relax=2)
    prot.n_stage = 2
    prot.skip_minimize = 1
    out = []
    monkeypatch.setenv("IGMK_SKIN", "0.7")                     # the two kernels default to different neighbour skins
    for force in (False, True, True):
        if force:
            monkeypatch.setenv("IGMK_FORCE_CLUSTER", str(cl))
        with MStepBatch(0) as b:
            slots = [b.submit(ps) for ps in systems]
            b.run(prot)
            out.append([b.fetch(s) for s in slots])
    for (a, ra), (c, rc), (e, _) in zip(*out):
        assert np.abs(a - c).max() <= 2e-3 and ra.n_rebuilds == rc.n_rebuilds
        assert np.array_equal(c.view(np.uint32), e.view(np.uint32))


@pytest.mark.parametrize("cl", [2, 8])
def test_cluster_md_kernel_force_loop_matches_oracle(monkeypatch, cl):
    """the forces the cluster kernel itself computes (remote gathers, remote inbox flags, mailboxes) against the oracle"""
    monkeypatch.setenv("IGMK_FORCE_CLUSTER", str(cl))
    from helpers import demo_structure
    ffg = np.load(__import__("os").path.join(__import__("conftest").GOLDEN, "ff_golden.npz"))
    systems = [demo_structure(ffg, k) for k in (1, 4)]
    prot = short_protocol()
    with MStepBatch(0) as b:
        slots = [b.submit(ps) for ps in systems]
        b.md_forces(prot, 2)
        for ps, slot in zip(systems, slots):
            f_ref, _ = ff.single_point(ps, prot.evfactor[2], prot.envfactor[2], 1.0)
            f = b.fetch_forces(slot)
            assert np.abs(f - f_ref).max() <= 2e-4 * np.abs(f_ref).max(), np.abs(f - f_ref).max() / np.abs(f_ref).max()


def test_global_memory_md_kernel_force_loop_matches_oracle(monkeypatch):
    """the forces k_md_big itself computes (state in global memory, every bond walked from both ends, full and candidate
    lists) against the oracle: 3010-atom demo structures, forced onto that kernel, and the 9500-bead system that needs it"""
    monkeypatch.setenv("IGMK_FORCE_BIG", "1")
    from helpers import demo_structure
    ffg = np.load(__import__("os").path.join(__import__("conftest").GOLDEN, "ff_golden.npz"))
    prot = short_protocol()
    for systems in ([demo_structure(ffg, k) for k in (2, 7)], [_big_system(seed=6)[2]]):
        with MStepBatch(0) as b:
            slots = [b.submit(ps) for ps in systems]
            b.md_forces(prot, 1)
            for ps, slot in zip(systems, slots):
                f_ref, _ = ff.single_point(ps, prot.evfactor[1], prot.envfactor[1], 1.0)
                f = b.fetch_forces(slot)
                assert np.abs(f - f_ref).max() <= 2e-4 * np.abs(f_ref).max(), np.abs(f - f_ref).max() / np.abs(f_ref).max()


def test_single_point_and_cg_beyond_shared_memory():
    g, rad, ps = _big_system()
    f_ref, e_ref = ff.single_point(ps)
    prot = short_protocol()
    prot.n_stage = 0
    prot.max_cg_iter = 12; prot.max_cg_eval = 40
    with MStepBatch(0) as b:
        slot = b.submit(ps)
        b.single_point()
        _, res = b.fetch(slot)
        f = b.fetch_forces(slot)
        assert res.final_energy == pytest.approx(e_ref["total"], rel=1e-10)
        assert np.abs(f - f_ref).max() <= 1e-10 * np.abs(f_ref).max()
        b.single_point(f32=True)
        _, res32 = b.fetch(slot)
        assert res32.final_energy == pytest.approx(e_ref["total"], rel=2e-4)
        b.run(prot)                                   # CG only
        xyz, res = b.fetch(slot)
    x_ref, r_ref = ff.optimize(ps, prot)
    assert (res.cg_iters, res.cg_evals, res.cg_stop) == (r_ref.cg_iters, r_ref.cg_evals, r_ref.cg_stop)
    assert res.final_energy == pytest.approx(r_ref.final_energy, rel=1e-8)


def test_structures_between_4096_and_8192_atoms_leave_the_one_cta_kernel():
    """5700 beads (a 1 Mb genome) do not fit one SM's shared memory in the thread-per-atom layout: a lone structure takes
    a thread-block cluster, a batch of many the global-memory kernel; both agree with the oracle's single point and
    bring the energy down"""
    g, rad, ps = _big_system(n_chain=6, chain=950, seed=8)
    assert 4096 < len(g) <= 8192
    f_ref, e0 = ff.single_point(ps)
    prot = short_protocol(md=120, relax=20)
    with MStepBatch(0) as b:
        slot = b.submit(ps)
        b.md_forces(prot, 1)
        f = b.fetch_forces(slot)
        f1, _ = ff.single_point(ps, prot.evfactor[1], prot.envfactor[1], 1.0)
        assert np.abs(f - f1).max() <= 2e-4 * np.abs(f1).max()
        b.run(prot)
        xyz, res = b.fetch(slot)
    assert res.status == 0 and np.isfinite(xyz).all() and res.final_energy < 0.05 * e0["total"]
    with MStepBatch(0) as b:                       # 40 structures: clusters of two CTAs
        for k in range(40):
            b.submit(ps)
        b.run(prot)
        many = [b.fetch(k) for k in range(40)]
    assert all(r.status == 0 and r.final_energy < 0.05 * e0["total"] for _, r in many)
    assert all(np.array_equal(many[0][0], x) for x, _ in many[1:])
    with MStepBatch(0) as b:                       # more structures than clusters fit the GPU: one CTA each, state in global memory
        for k in range(80):
            b.submit(ps)
        b.run(prot)
        many = [b.fetch(k) for k in (0, 41, 79)]
    assert all(r.status == 0 and r.final_energy < 0.05 * e0["total"] for _, r in many)
    assert all(np.array_equal(many[0][0], x) for x, _ in many[1:])


def test_big_structure_full_protocol():
    g, rad, ps = _big_system(seed=4)
    _, e0 = ff.single_point(ps)
    prot = short_protocol(md=150, relax=30)
    with MStepBatch(0) as b:
        slot = b.submit(ps)
        b.run(prot)
        xyz, res = b.fetch(slot)
        br, _ = b.violation_ratios(slot)
    assert res.status == 0 and np.isfinite(xyz).all()
    assert res.final_energy < 0.05 * e0["total"]
    assert res.final_energy == pytest.approx(res.pair_energy + res.bond_energy + res.env_energy[0], rel=1e-9)
    assert np.all(xyz[len(g)] == 0.0)
    # the energies reported are those of the returned coordinates (checked with the oracle)
    ps2 = synth.pack_structure(xyz[:len(g)], rad, g, ps.bond_i[ps.bond_i.size - 3000:], ps.bond_j[ps.bond_j.size - 3000:],
                               ps.bond_r0[ps.bond_r0.size - 3000:], nucleus_radius=6000.0)
    _, e1 = ff.single_point(ps2)
    assert e1["total"] == pytest.approx(res.final_energy, rel=1e-3, abs=1.0)


def test_config4_size_200kb_genome():
    """hg38 male at 200 kb: 29 838 beads (BASELINE config 4).  Single point vs the oracle at full size
    and a few hundred MD steps + CG through the global-memory path."""
    g = synth.genome_200kb()
    assert len(g) == 29838
    R = 5500.0
    rad = synth.bead_radii(len(g), R, 0.4)
    rng = np.random.default_rng(2)
    xyz = synth.territories(g, R, rng)
    hi = rng.integers(0, len(g), 20000)
    hj = (hi + rng.integers(1, 50, 20000)) % len(g)
    ps = synth.pack_structure(xyz, rad, g, hi, hj, np.full(20000, 4.0 * float(rad[0])), nucleus_radius=R, seed=9)
    f_ref, e_ref = ff.single_point(ps)
    prot = short_protocol(md=40, relax=10)
    prot.max_cg_iter = 5; prot.max_cg_eval = 15
    with MStepBatch(0) as b:
        slot = b.submit(ps)
        b.single_point()
        _, res = b.fetch(slot)
        f = b.fetch_forces(slot)
        assert res.final_energy == pytest.approx(e_ref["total"], rel=1e-10)
        assert np.abs(f - f_ref).max() <= 1e-10 * np.abs(f_ref).max()
        b.md_forces(prot, 0)                       # the cluster kernel's own force loop at full size (8 CTAs x 4096 positions)
        f32 = b.fetch_forces(slot)
        f0, _ = ff.single_point(ps, prot.evfactor[0], prot.envfactor[0], 1.0)
        assert np.abs(f32 - f0).max() <= 2e-4 * np.abs(f0).max()
        b.run(prot)
        xyz2, res = b.fetch(slot)
        md_ms, cg_ms, _ = b.last_timings()
    assert res.status == 0 and np.isfinite(xyz2).all() and res.final_energy < e_ref["total"]
    import os
    os.environ["IGMK_FORCE_BIG"] = "1"             # the same forces from the one-CTA global-memory kernel at full size
    try:
        with MStepBatch(0) as b:
            slot = b.submit(ps)
            b.md_forces(prot, 0)
            fb = b.fetch_forces(slot)
    finally:
        del os.environ["IGMK_FORCE_BIG"]
    assert np.abs(fb - f0).max() <= 2e-4 * np.abs(f0).max()
    print("200 kb structure: %.1f ms for %d MD force evaluations, CG %.1f ms for %d evaluations"
          % (md_ms, 4 * (40 + 10 + 2), cg_ms, res.cg_evals))
