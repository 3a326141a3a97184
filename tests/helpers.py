"""Shared builders for the tests (not a test module)."""
import numpy as np

from igm_b200 import synth
from igm_b200._abi import PackedSystem, KIND_BEAD, KIND_STATIC, KIND_CENTROID, BOND_UPPER


def demo_structure(ff_golden, k, n_extra_bonds=4000, seed=0, with_damid=True, with_centroid=True, evfactor=1.0,
                   damid_axes=(5500.0 * 0.95, 5000.0 * 0.95, 4500.0 * 0.95)):
    """One of the six demo structures (3008 beads) dressed with every term of the force field:
    polymer bonds, random upper/lower bonds of several (k, r0) types, the nuclear envelope with its
    static centre dummy, a DamID-style k<0 envelope on a subset, a SPRITE-style moving centroid."""
    g = synth.genome_2mb()
    xyz = ff_golden["coordinates"][:, k, :]
    rad = ff_golden["radii"]
    n = len(g)
    rng = np.random.default_rng(seed + 17 * k)
    pos = [xyz, np.zeros((1, 3), dtype=np.float32)]
    radii = [rad, np.zeros(1, dtype=np.float32)]
    kind = [np.full(n, KIND_BEAD, dtype=np.uint8), np.array([KIND_STATIC], dtype=np.uint8)]
    bi, bj, r0 = synth.polymer_bonds(g, rad)
    bk = np.ones(bi.size)
    bs = np.full(bi.size, BOND_UPPER, dtype=np.uint8)
    # random restraints: half between near beads (active lower bounds), half far (active upper bounds)
    ei = rng.integers(0, n, n_extra_bonds)
    ej = (ei + rng.integers(1, n - 1, n_extra_bonds)) % n
    er0 = rng.choice([2.0, 3.0, 5.0, 9.0], n_extra_bonds) * (rad[ei] + rad[ej]).astype(np.float64)
    ek = rng.choice([0.5, 1.0, 2.5], n_extra_bonds)
    es = rng.integers(0, 2, n_extra_bonds).astype(np.uint8)
    bi = np.concatenate([bi, ei]); bj = np.concatenate([bj, ej]); r0 = np.concatenate([r0, er0])
    bk = np.concatenate([bk, ek]); bs = np.concatenate([bs, es])
    n_atoms = n + 1
    if with_centroid:
        members = rng.choice(n, 12, replace=False)
        pos.append(xyz[members].mean(axis=0, keepdims=True).astype(np.float32))
        radii.append(np.zeros(1, dtype=np.float32))
        kind.append(np.array([KIND_CENTROID], dtype=np.uint8))
        bi = np.concatenate([bi, members]); bj = np.concatenate([bj, np.full(12, n_atoms)])
        r0 = np.concatenate([r0, np.full(12, 300.0)]); bk = np.concatenate([bk, np.full(12, 1.0)])
        bs = np.concatenate([bs, np.full(12, BOND_UPPER, dtype=np.uint8)])
        n_atoms += 1
    # bonds to the static dummy (radial FISH style, fish.py:136-159)
    fi = rng.choice(n, 40, replace=False)
    bi = np.concatenate([bi, fi]); bj = np.concatenate([bj, np.full(40, n)])
    r0 = np.concatenate([r0, rng.uniform(1000, 5000, 40)]); bk = np.concatenate([bk, np.full(40, 1.0)])
    bs = np.concatenate([bs, rng.integers(0, 2, 40).astype(np.uint8)])
    env_abck = [(5500.0, 5500.0, 5500.0, 1.0)]
    env_members = [np.arange(n)]
    if with_damid:
        env_abck.append(tuple(damid_axes) + (-0.3,))
        env_members.append(np.sort(rng.choice(n, 300, replace=False)))
    env_off = np.concatenate([[0], np.cumsum([len(m) for m in env_members])])
    return PackedSystem(np.vstack(pos), np.concatenate(radii), np.concatenate(kind), bi, bj, bk, r0, bs,
                        np.array(env_abck), env_off, np.concatenate(env_members), evfactor, seed=1 + k, struct_id=k)


def small_system(seed=0, n_struct=1, sizes=(14, 11, 8, 6), n_autosomes=3, R=1400.0, n_hic=40):
    """A toy genome small enough for the CPU oracle to run the whole protocol in seconds."""
    g = synth.Genome(list(sizes), n_autosomes=n_autosomes)
    rad = synth.bead_radii(len(g), R, 0.3)
    out = []
    for s in range(n_struct):
        rng = np.random.default_rng(1000 * seed + s)
        xyz = synth.territories(g, R, rng)
        hi = rng.integers(0, len(g), n_hic)
        hj = (hi + rng.integers(1, len(g) - 1, n_hic)) % len(g)
        out.append(synth.pack_structure(xyz, rad, g, hi, hj, np.full(n_hic, 4.0 * float(rad[0])),
                                        nucleus_radius=R, seed=11 + s, struct_id=s))
    return g, rad, out


def short_protocol(md=200, relax=50, **kw):
    from igm_b200 import _abi
    p = _abi.make_protocol(synth.demo_protocol_options(), **kw)
    for k in range(p.n_stage):
        p.mdsteps[k] = md
    p.relax_mdsteps = relax
    return p


def population_stats(systems, coords, tol=0.05):
    """Per-structure statistics of a population, evaluated by the SAME code (numpy + the C oracle) whichever arm produced
    the coordinates: log10(1 + E_total), log10(1 + E_bond), E_pair per bead, number of violated bonds (ratio > tol,
    forces.py:128-136, :176-184), radius of gyration.  -> dict name -> float64 [S]"""
    from oracle import ff
    out = {k: [] for k in ("log_e_total", "log_e_bond", "e_pair_per_bead", "n_violated", "rg")}
    for ps, xyz in zip(systems, coords):
        q = PackedSystem(xyz, ps.radius, ps.kind, ps.bond_i, ps.bond_j, ps.bond_k, ps.bond_r0, ps.bond_style,
                         ps.env_abck, ps.env_off, ps.env_members, ps.evfactor, seed=ps.seed, struct_id=ps.struct_id)
        _, e = ff.single_point(q)
        nb = int((ps.kind == KIND_BEAD).sum())
        d = np.linalg.norm(xyz[ps.bond_i].astype(np.float64) - xyz[ps.bond_j].astype(np.float64), axis=1)
        viol = np.where(ps.bond_style == BOND_UPPER, d - ps.bond_r0, ps.bond_r0 - d)
        ratio = np.where(ps.bond_r0 > 0, np.maximum(0.0, viol) / np.where(ps.bond_r0 > 0, ps.bond_r0, 1.0), 0.0)
        x = xyz[:nb].astype(np.float64)
        out["log_e_total"].append(np.log10(1.0 + e["total"]))
        out["log_e_bond"].append(np.log10(1.0 + e["bond"]))
        out["e_pair_per_bead"].append(e["pair"] / nb)
        out["n_violated"].append(float(np.count_nonzero(ratio > tol)))
        out["rg"].append(float(np.sqrt(((x - x.mean(axis=0)) ** 2).sum(axis=1).mean())))
    return {k: np.array(v, dtype=np.float64) for k, v in out.items()}


def assert_population_stats_match(a, b, nsig=4.5, floors=None):
    """Criterion 3 of north_star as a test: two populations (dicts of population_stats) agree when every statistic's
    means differ by less than nsig standard errors of the difference, sqrt(var_a / S_a + var_b / S_b) -- plus an
    absolute floor per statistic for quantities whose spread is ~0 (all structures satisfied)."""
    floors = floors or {}
    for k in a:
        ma, mb = a[k].mean(), b[k].mean()
        se = np.sqrt(a[k].var(ddof=1) / a[k].size + b[k].var(ddof=1) / b[k].size)
        lim = nsig * se + floors.get(k, 0.0)
        assert abs(ma - mb) <= lim, "%s: means %.5g vs %.5g differ by %.3g > %.3g (%.1f standard errors)" % (
            k, ma, mb, abs(ma - mb), lim, abs(ma - mb) / max(se, 1e-300))
