"""BASELINE configs[2] end to end on a reduced genome: Hi-C + lamina DamID + ellipsoidal envelope + FISH pairs and
radial probes (the restraint set of igm-config_all.json).  One A/M iteration on the GPU against the same iteration
with the CPU oracles doing every step:
  * A-steps (activation distances, DamID distances, FISH ranks / targets) and all assignments: bit-exact;
  * the restraints every structure ends up with: identical (sizes, double-precision single point to 1e-10, the MD
    kernel's own force loop to float32 round-off);
  * the optimised population: per-structure statistics agree within standard errors (criterion 3)."""
import numpy as np
import pytest

from igm_b200 import synth, iteration, restraints_pop as rp
from igm_b200._abi import PackedSystem, BOND_UPPER
from igm_b200.astep import unpack_bits
from igm_b200.kernel import MStepBatch
from oracle import ff, astep_oracle as ao, damid_oracle as do, fish_oracle as fo, restraints_oracle as ro
from helpers import short_protocol, population_stats, assert_population_stats_match

pytestmark = pytest.mark.gpu

R = 1400.0
SEMI = (1700.0, 1400.0, 1150.0)


def _workload(S=64, seed=5):
    g = synth.Genome([14, 11, 8, 6], n_autosomes=3)
    n = len(g)
    rad = synth.bead_radii(n, R, 0.3)
    crd = synth.population_territories(g, S, R=1100.0, seed=seed)
    rng = np.random.default_rng(seed)
    pi, pj, pw = synth.hic_pairs(g, 0.004, seed=3)
    hic = {"pairs": (pi, pj, pw, np.zeros(pi.size)), "contact_range": 2.0, "k": 1.0}
    loci = np.arange(g.n_hap, dtype=np.int32)
    damid = {"loci": loci, "pwish": np.clip(rng.normal(0.3, 0.25, loci.size), 0, 1), "plast": np.zeros(loci.size),
             "contact_range": 0.05, "shape": "ellipsoid", "nucleus_param": SEMI, "k": 0.6}
    dip = np.nonzero(np.diff(g.copy_off) == 2)[0]
    probes = rng.choice(dip, 6, replace=False).astype(np.int32)
    pa = rng.choice(dip, 16, replace=False).astype(np.int32).reshape(8, 2)
    # "experimental" distributions: sorted samples of the population's own distances + noise
    rlo, rhi, _, _ = fo.fish_ranks(crd, g.copy_index, probes)
    plo, phi, _, _ = fo.fish_ranks(crd, g.copy_index, pa[:, 0], pa[:, 1])
    noisy = lambda v: np.sort(np.abs(v.astype(np.float64) + rng.normal(0, 40.0, v.shape)) + 150.0, axis=1)
    fish = {"rtype": "rRpP", "tol": 60.0, "k": 1.5, "center": n + 2, "probes": probes, "pairs": pa,
            "radial_min": noisy(rlo), "radial_max": noisy(rhi), "pair_min": noisy(plo), "pair_max": noisy(phi)}
    tmpl = synth.pack_structure(np.zeros((n, 3), dtype=np.float32), rad, g, semiaxes=SEMI, seed=0, struct_id=0, n_extra_dummies=2)
    return g, rad, crd, hic, damid, fish, tmpl


def _oracle_iteration(g, rad, crd, hic, damid, fish, tmpl):
    """every A-step and the per-structure restraint build with the CPU restatements of the reference"""
    S = crd.shape[1]
    pi, pj, pw, pl = hic["pairs"]
    rows = ao.actdist_rows(pi, pj, pw, pl, crd, rad, g.chrom, g.copy_index, hic["contact_range"])
    hmask = ao.assign(crd, rows[0], rows[1], rows[2])
    drows = do.damid_rows(damid["loci"], damid["pwish"], damid["plast"], crd, rad, g.copy_index, damid["contact_range"],
                          damid["shape"], np.array(damid["nucleus_param"]))
    dmask = do.damid_assign(crd, rad, drows[0], drows[1], damid["contact_range"], damid["nucleus_param"])
    _, _, rmin, rmax = fo.fish_ranks(crd, g.copy_index, fish["probes"])
    _, _, pmin, pmax = fo.fish_ranks(crd, g.copy_index, fish["pairs"][:, 0], fish["pairs"][:, 1])
    tg = {"radial_min": np.take_along_axis(fish["radial_min"], rmin.astype(np.int64), axis=1),
          "radial_max": np.take_along_axis(fish["radial_max"], rmax.astype(np.int64), axis=1),
          "pair_min": np.take_along_axis(fish["pair_min"], pmin.astype(np.int64), axis=1),
          "pair_max": np.take_along_axis(fish["pair_max"], pmax.astype(np.int64), axis=1)}
    abc = np.array(damid["nucleus_param"]) * (1 - damid["contact_range"])
    systems = []
    for s in range(S):
        ff_ = ro.fish_forces(crd[:, s, :], g.copy_index, fish["center"], fish["rtype"], fish["tol"], fish["k"], s,
                             probes=fish["probes"], pairs=fish["pairs"], **tg)
        m = hmask[:, s]
        r0 = (hic["contact_range"] * (rad[rows[0][m]] + rad[rows[1][m]])).astype(np.float64)
        bi = np.concatenate([tmpl.bond_i, rows[0][m], [f[1] for f in ff_]]).astype(np.int32)
        bj = np.concatenate([tmpl.bond_j, rows[1][m], [f[2] for f in ff_]]).astype(np.int32)
        bk = np.concatenate([tmpl.bond_k, np.full(int(m.sum()), hic["k"]), [f[4] for f in ff_]])
        br = np.concatenate([tmpl.bond_r0, r0, [f[3] for f in ff_]])
        bs = np.concatenate([tmpl.bond_style, np.full(int(m.sum()), BOND_UPPER), [f[0] for f in ff_]]).astype(np.uint8)
        mem = drows[0][dmask[:, s]]
        xyz = tmpl.xyz.copy()
        xyz[:len(g)] = crd[:, s, :]
        systems.append(PackedSystem(xyz, tmpl.radius, tmpl.kind, bi, bj, bk, br, bs,
                                    np.vstack([tmpl.env_abck, [abc[0], abc[1], abc[2], -damid["k"]]]),
                                    np.concatenate([tmpl.env_off, [tmpl.env_off[-1] + mem.size]]),
                                    np.concatenate([tmpl.env_members, mem]), tmpl.evfactor, seed=11 + s, struct_id=s))
    return rows, hmask, drows, dmask, tg, systems


def test_config3_iteration_matches_the_cpu_oracles():
    g, rad, crd, hic, damid, fish, tmpl = _workload()
    S, n = crd.shape[1], len(g)
    rows, hmask, drows, dmask, tg, systems = _oracle_iteration(g, rad, crd, hic, damid, fish, tmpl)
    assert hmask.sum() > 10 * S and 0 < dmask.sum() < dmask.size

    # ---- A-steps + assignments: bit-exact ----
    asm = iteration.assemble_restraints(crd, g, rad, hic=hic, damid=damid, fish=fish)
    for a, b in zip(asm["hic_rows"], rows):
        assert np.array_equal(a, b)
    for a, b in zip(asm["damid_rows"], drows):
        assert np.array_equal(a, b)
    assert np.array_equal(unpack_bits(asm["hic_bits"], S), hmask) and np.array_equal(unpack_bits(asm["damid_bits"], S), dmask)
    for k in tg:
        assert np.array_equal(asm["fish_targets"][k], tg[k])

    # ---- the restraints of every structure: identical ----
    row, col, r0, k_hic, bits, extras = iteration.population_inputs(asm, rad, S, hic=hic, damid=damid)
    prot = short_protocol(md=250, relax=40)
    seeds = np.arange(S, dtype=np.int64) + 11
    with MStepBatch(0) as b:
        b.submit_population(tmpl, crd, n, row, col, r0, k_hic, bits, seeds=seeds, extras=extras)
        for s in range(S):
            assert b.slot_sizes(s) == (n + 3, systems[s].bond_i.size, systems[s].env_members.size)
        b.single_point()
        ref = [ff.single_point(ps) for ps in systems]
        for s in range(S):
            _, res = b.fetch(s)
            f_ref, e_ref = ref[s]
            assert res.final_energy == pytest.approx(e_ref["total"], rel=1e-10)
            assert res.bond_energy == pytest.approx(e_ref["bond"], rel=1e-10, abs=1e-10 * e_ref["total"])
            assert res.env_energy[1] == pytest.approx(e_ref["env"][1], rel=1e-10, abs=1e-10 * e_ref["total"])
            assert np.abs(b.fetch_forces(s) - f_ref).max() <= 1e-10 * np.abs(f_ref).max()
        # the MD kernel's own force loop on this restraint mix (lower bounds, many bond types, a k < 0 envelope)
        b.md_forces(prot, 1)
        for s in range(0, S, 7):
            f_ref, _ = ff.single_point(systems[s], prot.evfactor[1], prot.envfactor[1], 1.0)
            assert np.abs(b.fetch_forces(s) - f_ref).max() <= 2e-4 * np.abs(f_ref).max()

    # ---- one full iteration: the optimised populations agree statistically (criterion 3) ----
    out = iteration.am_iteration(crd, g, rad, tmpl, synth.demo_protocol_options() | {"custom_annealing_protocol": _short_cap(250, 40)},
                                 hic=hic, damid=damid, fish=fish, seeds=seeds, violations=True)
    gpu_xyz = []
    for s in range(S):
        x = systems[s].xyz.copy()
        x[:n] = out["coordinates"][:, s, :]
        gpu_xyz.append(x)
    cpu_xyz = [ff.optimize(ps, prot)[0].astype(np.float32) for ps in systems]
    st_gpu, st_cpu = population_stats(systems, gpu_xyz), population_stats(systems, cpu_xyz)
    assert_population_stats_match(st_gpu, st_cpu, floors={"n_violated": 0.25, "log_e_total": 0.02, "log_e_bond": 0.02,
                                                          "e_pair_per_bead": 1e-3, "rg": 1.0})
    # the optimisation did something: far fewer violated restraints than at the start
    st0 = population_stats(systems, [ps.xyz for ps in systems])
    assert st_gpu["n_violated"].mean() < 0.5 * st0["n_violated"].mean()
    assert all(v[0] == ps.bond_i.size for v, ps in zip(out["violations"], systems))


def _short_cap(md, relax):
    cap = dict(synth.demo_protocol_options()["custom_annealing_protocol"])
    cap["mdsteps"] = [md] * 4
    cap["relax"] = dict(cap["relax"], mdsteps=relax)
    return cap


def test_extras_are_sliced_per_chunk_like_the_bit_matrix():
    """run_population submits in chunks of 32: the per-structure bonds and envelope bits must follow the chunk"""
    g, rad, crd, hic, damid, fish, tmpl = _workload(S=70, seed=9)
    S, n = crd.shape[1], len(g)
    asm = iteration.assemble_restraints(crd, g, rad, hic=hic, damid=damid, fish=fish)
    row, col, r0, k_hic, bits, extras = iteration.population_inputs(asm, rad, S, hic=hic, damid=damid)
    with MStepBatch(0) as a, MStepBatch(0) as b:
        a.submit_population(tmpl, crd, n, row, col, r0, k_hic, bits, extras=extras)
        for lo in range(0, S, 32):
            hi = min(S, lo + 32)
            b.submit_population(tmpl, crd, n, row, col, r0, k_hic, bits, s_begin=lo, n_struct=hi - lo, bits_word0=lo // 32, extras=extras)
        a.single_point(); b.single_point()
        for s in range(S):
            assert a.slot_sizes(s) == b.slot_sizes(s) and a.fetch(s)[1].final_energy == b.fetch(s)[1].final_energy
    with pytest.raises(RuntimeError, match="negative bound distance"):
        bad = dict(fish, tol=1e6)
        asm2 = iteration.assemble_restraints(crd, g, rad, fish=bad)
        with MStepBatch(0) as c:
            c.submit_population(tmpl, crd, n, row, col, r0, k_hic, bits, extras=rp.extras(S, bonds=asm2["fish_bonds"]))
