"""M-step protocol on the GPU: annealing MD (float32) + CG (float64) against the CPU oracle.
Trajectories are chaotic, so agreement is statistical (north_star criterion 3) except where the
arithmetic is deterministic: a run repeated on the GPU must reproduce itself bit for bit."""
import numpy as np
import pytest

from igm_b200 import synth
from igm_b200.kernel import MStepBatch, optimize, optimize_population
from oracle import ff
from helpers import small_system, short_protocol, demo_structure

pytestmark = pytest.mark.gpu


def _violations(ps, xyz, tol=0.05):
    """HarmonicUpperBound.getViolationRatio (forces.py:128-136) over the structure's bonds"""
    d = np.linalg.norm(xyz[ps.bond_i] - xyz[ps.bond_j], axis=1)
    ratio = np.where(ps.bond_r0 > 0, np.maximum(0.0, d - ps.bond_r0) / np.maximum(ps.bond_r0, 1e-30), 0.0)
    return int((ratio > tol).sum())


def test_cg_only_matches_oracle_minimiser():
    """no MD stages: the double-precision CG alone is deterministic and follows the oracle's
    path step for step (same iteration / evaluation counts, energies to 1e-9)"""
    g, rad, systems = small_system(seed=4, n_struct=3)
    prot = short_protocol()
    prot.n_stage = 0
    with MStepBatch(0) as b:
        slots = [b.submit(ps) for ps in systems]
        b.run(prot)
        for ps, slot in zip(systems, slots):
            x_ref, r_ref = ff.optimize(ps, prot)
            xyz, res = b.fetch(slot)
            assert (res.cg_iters, res.cg_evals, res.cg_stop) == (r_ref.cg_iters, r_ref.cg_evals, r_ref.cg_stop)
            assert res.final_energy == pytest.approx(r_ref.final_energy, rel=1e-8, abs=1e-10)
            assert np.abs(xyz - x_ref).max() <= 1e-3      # float32 write-back of ~1e3-sized coordinates


def test_full_protocol_statistics_match_oracle():
    """criterion 3 on the M-step alone: 64 structures through the GPU protocol and through the CPU oracle (independent
    arithmetic: float32 vs float64, so the chaotic trajectories differ) end in the same population -- per-structure
    energies, violated restraints and radius of gyration agree within 4.5 standard errors of the means"""
    from helpers import population_stats, assert_population_stats_match
    S = 64
    g, rad, systems = small_system(seed=5, n_struct=S)
    prot = short_protocol(md=600, relax=100)
    gpu_xyz, cpu_xyz, t_gpu = [], [], []
    with MStepBatch(0) as b:
        slots = [b.submit(ps) for ps in systems]
        b.run(prot)
        for ps, slot in zip(systems, slots):
            xyz, res = b.fetch(slot)
            assert res.status == 0 and np.isfinite(xyz).all()
            assert res.final_energy == pytest.approx(res.pair_energy + res.bond_energy + res.env_energy[0], rel=1e-9, abs=1e-9)
            assert np.all(xyz[len(g)] == 0.0)                          # static dummy untouched
            assert (np.linalg.norm(xyz[:len(g)], axis=1) + rad <= 1400.0 * 1.02).all()
            gpu_xyz.append(xyz); t_gpu.append(res.temperature)
            cpu_xyz.append(ff.optimize(ps, prot)[0].astype(np.float32))
    st_gpu, st_cpu = population_stats(systems, gpu_xyz), population_stats(systems, cpu_xyz)
    assert_population_stats_match(st_gpu, st_cpu, floors={"n_violated": 0.1, "log_e_total": 0.01, "log_e_bond": 0.01,
                                                          "e_pair_per_bead": 1e-4, "rg": 0.5})
    n_restr = sum(ps.bond_i.size for ps in systems)
    assert st_gpu["n_violated"].sum() <= 0.01 * n_restr and st_cpu["n_violated"].sum() <= 0.01 * n_restr
    assert max(t_gpu) < 0.2                                            # annealed to T -> 0


def test_wrong_thermostat_or_prefactor_would_be_caught():
    """the statistical comparison has teeth: the same GPU run with the bond prefactor halved (E = k/2 dr^2), or cut off
    after the hottest stage, leaves the population measurably different from the oracle's"""
    from helpers import population_stats, assert_population_stats_match
    S = 64
    g, rad, systems = small_system(seed=5, n_struct=S, n_hic=120)
    prot = short_protocol(md=600, relax=100)
    prot.skip_minimize = 1                         # the minimiser would hide what the dynamics did
    cpu_xyz = [ff.optimize(ps, prot)[0].astype(np.float32) for ps in systems]
    st_cpu = population_stats(systems, cpu_xyz)
    floors = {"n_violated": 0.1, "log_e_total": 0.01, "log_e_bond": 0.01, "e_pair_per_bead": 1e-4, "rg": 0.5}

    def gpu(p):
        with MStepBatch(0) as b:
            slots = [b.submit(ps) for ps in systems]
            b.run(p)
            return population_stats(systems, [b.fetch(s)[0] for s in slots])

    assert_population_stats_match(gpu(prot), st_cpu, floors=floors)           # the right protocol passes ...
    hot = short_protocol(md=600, relax=100)
    hot.skip_minimize = 1
    hot.n_stage = 1                                                            # ... stopping at 500 K does not
    with pytest.raises(AssertionError):
        assert_population_stats_match(gpu(hot), st_cpu, floors=floors)


def test_run_is_bitwise_reproducible():
    g, rad, systems = small_system(seed=6, n_struct=4)
    prot = short_protocol(md=300, relax=50)
    out = []
    for rep in range(2):
        with MStepBatch(0) as b:
            slots = [b.submit(ps) for ps in systems]
            b.run(prot)
            out.append([b.fetch(s)[0] for s in slots])
    for a, c in zip(*out):
        assert np.array_equal(a.view(np.uint32), c.view(np.uint32))


def test_reset_coordinates_repeats_the_run():
    g, rad, systems = small_system(seed=7, n_struct=2)
    prot = short_protocol(md=100, relax=20)
    with MStepBatch(0) as b:
        slots = [b.submit(ps) for ps in systems]
        b.run(prot)
        first = [b.fetch(s)[0] for s in slots]
        b.reset_coordinates()
        b.run(prot)
        second = [b.fetch(s)[0] for s in slots]
    for a, c in zip(first, second):
        assert np.array_equal(a, c)


def test_thermostat_holds_the_target_temperature():
    """skip_minimize + a flat T=50 stage: the thermo temperature at the end is the target
    (temp/rescale window 0.1, fraction 1)"""
    g, rad, systems = small_system(seed=8)
    prot = short_protocol(md=300, relax=0)
    prot.n_stage = 1
    prot.tstart[0] = 50.0; prot.tstop[0] = 50.0
    prot.skip_minimize = 1
    with MStepBatch(0) as b:
        slot = b.submit(systems[0])
        b.run(prot)
        _, res = b.fetch(slot)
    n = systems[0].n_atoms
    want = 50.0 * (3.0 * (n - 1) - 3.0) / (3.0 * n - 3.0)    # thermo `temp` is over group all
    assert res.temperature == pytest.approx(want, abs=0.11)


def test_demo_size_structure_runs(ff_golden):
    """a 3008-bead structure with ~7k bonds through a shortened protocol"""
    ps = demo_structure(ff_golden, 0, n_extra_bonds=0, with_damid=False, with_centroid=False)
    prot = short_protocol(md=200, relax=50)
    with MStepBatch(0) as b:
        slot = b.submit(ps)
        b.run(prot)
        xyz, res = b.fetch(slot)
    assert res.status == 0 and np.isfinite(res.final_energy)
    assert res.final_energy < 168097.0                      # lower than where it started
    assert res.n_rebuilds > 1


class _P(object):
    def __init__(self, pos, r, ptype):
        self.pos = np.array(pos, dtype=np.float32); self.r = np.float32(r); self.ptype = ptype


class _F(object):
    def __init__(self, ftype, **kw):
        self.ftype = ftype
        self.__dict__.update(kw)


class _Model(object):
    def __init__(self, uid):
        self.id = uid; self.particles = []; self.forces = []


def _toy_model(uid, seed):
    """what ModelingStep.task builds (ModelingStep.py:184-264), with duck-typed igm classes"""
    g, rad, systems = small_system(seed=seed)
    ps = systems[0]
    m = _Model(uid)
    n = len(g)
    for i in range(n):
        m.particles.append(_P(ps.xyz[i], rad[i], 0))
    m.forces.append(_F(0, k=1.0))                                              # Steric
    m.particles.append(_P([0, 0, 0], 0.0, 1))                                  # Envelope's dummy
    m.forces.append(_F(3, shape="ellipsoid", semiaxes=np.array([1400.0] * 3), particle_ids=list(range(n)), k=1.0))
    for i, j, d, k in zip(ps.bond_i, ps.bond_j, ps.bond_r0, ps.bond_k):
        m.forces.append(_F(1, i=int(i), j=int(j), d=float(d), k=float(k)))
    return m, ps


def test_optimize_plugin_contract():
    """igm.model.kernel.lammps.optimize's contract (lammps.py:361-489): in-place positions, info keys"""
    m, ps = _toy_model(3, seed=9)
    opts = synth.demo_protocol_options()
    opts["custom_annealing_protocol"]["mdsteps"] = [100, 100, 100, 100]
    opts["custom_annealing_protocol"]["relax"]["mdsteps"] = 20
    cfg = {"optimization": {"kernel": "b200", "optimizer_options": opts, "kernel_opts": {"b200": {"seed": 6535}}},
           "runtime": {"step_no": 3}}
    before = np.array([p.pos for p in m.particles])
    info = optimize(m, cfg)
    after = np.array([p.pos for p in m.particles])
    assert set(info) == {"final-energy", "pair-energy", "bond-energy", "thermo", "md-time"}
    assert set(info["thermo"]) == {"Temp", "E_pair", "E_bond", "f_envelope0"}
    assert after.dtype == np.float32 and after.shape == before.shape and not np.array_equal(after, before)
    assert np.all(after[-1] == 0.0)
    assert info["final-energy"] == pytest.approx(info["pair-energy"] + info["bond-energy"] + info["thermo"]["f_envelope0"], abs=1e-6)
    assert info["md-time"] > 0
    # a population call gives the same answer for the same model / seed as the single call
    m2, _ = _toy_model(3, seed=9)
    m3, _ = _toy_model(4, seed=10)
    infos = optimize_population([m2, m3], cfg)
    assert np.array_equal(np.array([p.pos for p in m2.particles]), after)
    assert infos[0]["final-energy"] == info["final-energy"]


def test_optimize_raises_runtime_error_with_model_id():
    m, _ = _toy_model(12, seed=11)
    m.forces.append(_F(1, i=0, j=10 ** 6, d=1.0, k=1.0))
    cfg = {"optimization": {"kernel": "b200", "optimizer_options": synth.demo_protocol_options(), "kernel_opts": {}},
           "runtime": {"step_no": 0}}
    with pytest.raises(RuntimeError, match="modelid: 12"):
        optimize(m, cfg)


def test_submit_population_equals_per_structure_submits():
    """igmk_submit_population (template + slab + assignment bits) builds the same batch as one
    igmk_submit per structure: identical single points and bit-identical optimised coordinates"""
    from igm_b200.astep import Population
    from oracle import astep_oracle as ao
    g = synth.Genome([14, 11, 8, 6], n_autosomes=3)
    R = 1400.0
    rad = synth.bead_radii(len(g), R, 0.3)
    S = 37
    crd = synth.population_territories(g, S, R=R, seed=12)
    pi, pj, pw = synth.hic_pairs(g, 0.004, seed=3)
    with Population(S, rad, g.chrom, g.copy_off, g.copy_beads) as pop:
        pop.upload(crd)
        row, col, dist, _ = pop.actdist(pi, pj, pw, np.zeros(pi.size), 2.0)
        bits = pop.assign(row, col, dist)
    mask = ao.assign(crd, row, col, dist)
    r0 = (2.0 * (rad[row] + rad[col])).astype(np.float64)
    tmpl = synth.pack_structure(np.zeros((len(g), 3), dtype=np.float32), rad, g, nucleus_radius=R, seed=0, struct_id=100)
    seeds = np.arange(S, dtype=np.int64) + 5
    prot = short_protocol(md=150, relax=30)
    with MStepBatch(0) as a, MStepBatch(0) as b:
        first = a.submit_population(tmpl, crd, len(g), row, col, r0, 1.0, bits, seeds=seeds)
        assert first == 0 and len(a) == S
        for s in range(S):
            m = mask[:, s]
            b.submit(synth.pack_structure(crd[:, s, :], rad, g, row[m], col[m], r0[m], nucleus_radius=R, seed=int(seeds[s]),
                                          struct_id=100 + s))
        assert all(a.slot_sizes(s) == b.slot_sizes(s) for s in range(S))
        assert mask.sum() > 50 and sum(a.slot_sizes(s)[1] for s in range(S)) == S * tmpl.bond_i.size + int(mask.sum())
        a.single_point(); b.single_point()
        e_sp = [b.fetch(s)[1].final_energy for s in range(S)]
        for s in range(S):
            assert a.fetch(s)[1].final_energy == e_sp[s]
        a.run(prot); b.run(prot)
        for s in range(S):
            xa, ra = a.fetch(s); xb, rb = b.fetch(s)
            assert np.array_equal(xa.view(np.uint32), xb.view(np.uint32)) and ra.final_energy == rb.final_energy
        # a sub-range of the population with its own bit matrix
        with Population(10, rad, g.chrom, g.copy_off, g.copy_beads) as pop2, MStepBatch(0) as c:
            sub = np.ascontiguousarray(crd[:, 20:30, :])
            pop2.upload(sub)
            bits2 = pop2.assign(row, col, dist)
            c.submit_population(tmpl, crd, len(g), row, col, r0, 1.0, bits2, s_begin=20, n_struct=10, seeds=seeds[20:30])
            c.single_point()
            for k in range(10):
                assert c.fetch(k)[1].final_energy == e_sp[20 + k]


def test_population_without_seed_list_gives_every_structure_its_own_seed():
    """seeds=None: structure s runs `velocity create` with ((max(tmpl.seed, 1) * (tmpl.struct_id + s)) mod 9190037) + 1
    (lammps.py:159-165), not one shared stream -- identical start coordinates must end in different places, and exactly
    where an explicit list of those seeds puts them"""
    g = synth.Genome([14, 11, 8, 6], n_autosomes=3)
    R = 1400.0
    rad = synth.bead_radii(len(g), R, 0.3)
    one = synth.territories(g, R, np.random.default_rng(3))
    S = 4
    crd = np.ascontiguousarray(np.repeat(one[:, None, :], S, axis=1))
    none_i = np.zeros(0, dtype=np.int32)
    bits = np.zeros((0, 1), dtype=np.uint32)
    prot = short_protocol(md=120, relax=20)
    out = []
    for tmpl_seed, seeds in ((0, None), (0, np.array([((1 * (50 + s)) % 9190037) + 1 for s in range(S)], dtype=np.int64)),
                             (977, None), (977, np.array([((977 * (50 + s)) % 9190037) + 1 for s in range(S)], dtype=np.int64))):
        tmpl = synth.pack_structure(np.zeros((len(g), 3), dtype=np.float32), rad, g, nucleus_radius=R, seed=tmpl_seed, struct_id=50)
        with MStepBatch(0) as b:
            b.submit_population(tmpl, crd, len(g), none_i, none_i, np.zeros(0), 1.0, bits, seeds=seeds)
            b.run(prot)
            out.append([b.fetch(s)[0] for s in range(S)])
    for k in (0, 2):
        for s in range(S):
            assert np.array_equal(out[k][s].view(np.uint32), out[k + 1][s].view(np.uint32))
            for t in range(s):
                assert not np.array_equal(out[k][s], out[k][t])
    assert not np.array_equal(out[0][1], out[2][1])


def test_run_population_pipeline_equals_one_batch():
    """kernel.run_population (chunks of 32 on their own streams, assembly overlapping the MD) returns
    bit-identical coordinates, energies and violation counts to one batch holding every structure"""
    from igm_b200.astep import Population
    from igm_b200.kernel import run_population
    g = synth.Genome([14, 11, 8, 6], n_autosomes=3)
    R = 1400.0
    rad = synth.bead_radii(len(g), R, 0.3)
    S = 75                                            # 32 + 32 + ragged 11
    crd = synth.population_territories(g, S, R=R, seed=21)
    pi, pj, pw = synth.hic_pairs(g, 0.004, seed=4)
    with Population(S, rad, g.chrom, g.copy_off, g.copy_beads) as pop:
        pop.upload(crd)
        row, col, dist, _ = pop.actdist(pi, pj, pw, np.zeros(pi.size), 2.0)
        bits = pop.assign(row, col, dist)
    r0 = (2.0 * (rad[row] + rad[col])).astype(np.float64)
    tmpl = synth.pack_structure(np.zeros((len(g), 3), dtype=np.float32), rad, g, nucleus_radius=R, seed=0, struct_id=0)
    seeds = np.arange(S, dtype=np.int64) + 9
    prot = short_protocol(md=150, relax=30)
    new, results, viol, tim = run_population(tmpl, crd, len(g), row, col, r0, 1.0, bits, prot, seeds=seeds, chunk=32,
                                             violations_tol=0.05)
    assert tim["chunks"] == 3 and new.shape == crd.shape and len(results) == S
    with MStepBatch(0) as a:
        a.submit_population(tmpl, crd, len(g), row, col, r0, 1.0, bits, seeds=seeds)
        a.run(prot)
        for s in range(S):
            xa, ra = a.fetch(s)
            assert np.array_equal(xa[:len(g)].view(np.uint32), new[:, s, :].view(np.uint32))
            assert ra.final_energy == results[s].final_energy and results[s].status == 0
            br, _ = a.violation_ratios(s)
            assert viol[s][:2] == (br.size, int(np.count_nonzero(br > 0.05))) and viol[s][2] == len(g) and viol[s][3] <= 2
        slab = np.zeros((len(g), S + 5, 3), dtype=np.float32)              # bulk fetch into a wider slab at an offset
        res = a.fetch_population(slab, s_begin=5)
        assert np.array_equal(slab[:, 5:, :].view(np.uint32), new.view(np.uint32)) and not slab[:, :5, :].any()
        assert [r.final_energy for r in res] == [r.final_energy for r in results]
        with pytest.raises(RuntimeError):
            a.fetch_population(slab, s_begin=6)
    assert not np.array_equal(new, crd)
    with pytest.raises(ValueError):
        with MStepBatch(0) as b:
            b.submit_population(tmpl, crd, len(g), row, col, r0, 1.0, bits, s_begin=64, n_struct=11, bits_word0=3)


def test_config3_restraint_set_through_md(ff_golden):
    """2 Mb structure with every restraint kind of BASELINE config 3 at once (Hi-C-like upper bounds, lower
    bounds, an ellipsoidal k<0 lamina envelope, bonds to the static centre as radial FISH uses them, a
    SPRITE centroid): the MD + CG path runs it, lowers the energy, repeats itself bit for bit, and its
    reported energies are those of the returned coordinates."""
    ps = demo_structure(ff_golden, 8, n_extra_bonds=3000)
    prot = short_protocol(md=120, relax=30)
    outs = []
    for rep in range(2):
        with MStepBatch(0) as b:
            slot = b.submit(ps)
            b.single_point()
            e0 = b.fetch(slot)[1].final_energy
            b.reset_coordinates()
            b.run(prot)
            outs.append(b.fetch(slot))
    (xa, ra), (xb, rb) = outs
    assert np.array_equal(xa.view(np.uint32), xb.view(np.uint32)) and ra.final_energy == rb.final_energy
    assert ra.status == 0 and ra.n_env == 2 and ra.final_energy < e0
    assert np.all(xa[3008] == 0.0) and not np.array_equal(xa[3009], ps.xyz[3009])      # dummy fixed, centroid moved
    ps.xyz[:] = xa
    _, e = ff.single_point(ps)
    assert e["total"] == pytest.approx(ra.final_energy, rel=1e-4)
    assert e["env"][1] == pytest.approx(ra.env_energy[1], rel=1e-3, abs=1e-3 * abs(e["total"]))


def test_controller_batches_real_optimize_calls():
    """ModelingStep-style tasks under B200Controller: one launch, same results as optimize_population"""
    from igm_b200.controller import B200Controller
    opts = synth.demo_protocol_options()
    opts["custom_annealing_protocol"]["mdsteps"] = [60, 60, 60, 60]
    opts["custom_annealing_protocol"]["relax"]["mdsteps"] = 10
    cfg = {"optimization": {"kernel": "b200", "optimizer_options": opts, "kernel_opts": {"b200": {"seed": 77}}},
           "runtime": {"step_no": 1}}

    def task(i):
        m, _ = _toy_model(i + 1, seed=20 + i)
        info = optimize(m, cfg)
        return np.array([p.pos for p in m.particles]), info["final-energy"]

    ctl = B200Controller(max_threads=16)
    out = ctl.map(task, range(6))
    assert ctl.last_batcher.batches == [6]
    models = [_toy_model(i + 1, seed=20 + i)[0] for i in range(6)]
    infos = optimize_population(models, cfg)
    for (xyz, e), m, info in zip(out, models, infos):
        assert np.array_equal(xyz, np.array([p.pos for p in m.particles])) and e == info["final-energy"]


@pytest.mark.parametrize("uniform", [True, False])
def test_short_trajectory_matches_oracle(uniform):
    """same Park-Miller velocity draws, same integrator: after a few steps of the first stage the float32
    GPU trajectory still sits on the float64 oracle trajectory (both radius code paths of the MD kernel)"""
    g, rad, systems = small_system(seed=51, n_struct=2)
    if not uniform:
        for ps in systems:
            ps.radius[:len(g):3] *= np.float32(0.8)
            ps.radius[1:len(g):5] *= np.float32(1.15)
    prot = short_protocol(md=6, relax=0)
    prot.n_stage = 1
    prot.skip_minimize = 1
    with MStepBatch(0) as b:
        slots = [b.submit(ps) for ps in systems]
        b.run(prot)
        for ps, slot in zip(systems, slots):
            xyz, res = b.fetch(slot)
            x_ref, r_ref = ff.optimize(ps, prot)
            moved = np.abs(x_ref - ps.xyz).max()
            assert moved > 50.0                                     # T = 5000: the beads really travelled
            assert np.abs(xyz - x_ref).max() <= 2e-3 * moved
            assert res.temperature == pytest.approx(r_ref.temperature, rel=1e-3)


def test_optimize_population_over_all_devices():
    """kernel_opts/b200/devices = "all": the same results as on one device, whatever the split
    (one GPU on the test box: the code path still runs; several: contiguous blocks per device)"""
    opts = synth.demo_protocol_options()
    opts["custom_annealing_protocol"]["mdsteps"] = [60, 60, 60, 60]
    opts["custom_annealing_protocol"]["relax"]["mdsteps"] = 10
    base = {"optimizer_options": opts, "kernel": "b200"}
    cfg1 = {"optimization": dict(base, kernel_opts={"b200": {"seed": 5}}), "runtime": {"step_no": 0}}
    cfgn = {"optimization": dict(base, kernel_opts={"b200": {"seed": 5, "devices": "all"}}), "runtime": {"step_no": 0}}
    cfg2 = {"optimization": dict(base, kernel_opts={"b200": {"seed": 5, "devices": [0, 0, 0]}}), "runtime": {"step_no": 0}}
    ref = [_toy_model(i + 1, seed=60 + i)[0] for i in range(7)]
    want = optimize_population(ref, cfg1)
    for cfg in (cfgn, cfg2):
        ms = [_toy_model(i + 1, seed=60 + i)[0] for i in range(7)]
        got = optimize_population(ms, cfg)
        assert [g["final-energy"] for g in got] == [w["final-energy"] for w in want]
        for a, b in zip(ms, ref):
            assert np.array_equal(np.array([p.pos for p in a.particles]), np.array([p.pos for p in b.particles]))
