#!/usr/bin/env python
"""BASELINE configs[2] and configs[3] at their real genome sizes, one A/M iteration each on one GPU, as JSON lines
(bench.py embeds them under "extra_configs" at N = 1; `python tools/bench_configs.py [c3|c4]` prints them alone).

c3: 2 Mb diploid (3008 beads), Hi-C + lamina DamID + ellipsoidal envelope (7840, 6470, 2450) + FISH: 500 pairs and
    300 radial probes, rtype "rRpP" (restraint set of igm-config_all.json)
c4: 200 kb diploid (29 838 beads), Hi-C + SPRITE, through the cluster-per-structure MD kernel (a thread-block cluster
    per structure: 16 CTAs up to 7 structures, 8 CTAs in waves of 15 beyond -- upload() in igmk.cu)
Both are run on FEWER structures than the 1000 of the BASELINE text so that the default bench stays within minutes;
the number is in the line.  Synthetic inputs, seeded."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from igm_b200 import synth, iteration, sprite, restraints_pop as rp, _abi        # noqa: E402
from igm_b200.astep import Population                                             # noqa: E402
from igm_b200.kernel import run_population, merge_bond_lists                      # noqa: E402

SEMI = (7840.0, 6470.0, 2450.0)


def _freq_pairs(pop, g, n_pairs):
    iu, ju = np.triu_indices(g.n_hap, 1)
    freq = pop.contact_freq(iu.astype(np.int32), ju.astype(np.int32), 2.0)
    pw = np.minimum(np.float32(1.0), np.float32(1.1) * freq).astype(np.float32)
    thr = np.partition(pw, pw.size - n_pairs)[pw.size - n_pairs]
    keep = pw >= thr
    return (iu[keep].astype(np.int32), ju[keep].astype(np.int32), pw[keep].astype(np.float64),
            np.round(freq[keep].astype(np.float64), 4).astype(np.float32).astype(np.float64))


def run_c3(device=0, S=296, n_pairs=300000):
    g = synth.genome_2mb()
    n = len(g)
    rad = np.full(n, (0.4 * SEMI[0] * SEMI[1] * SEMI[2] / n) ** (1.0 / 3.0), dtype=np.float32)     # occupancy 0.4 of the ellipsoid
    opts = synth.demo_protocol_options()
    rng = np.random.default_rng(11)
    start = synth.population_territories(g, S, R=5500.0, seed=4321) * (np.array(SEMI, dtype=np.float32) / np.float32(5500.0) * np.float32(0.9))
    tmpl = synth.pack_structure(np.zeros((n, 3), dtype=np.float32), rad, g, semiaxes=SEMI, seed=0, struct_id=0, n_extra_dummies=2)
    crd, _ = iteration.relax_init(np.ascontiguousarray(start, dtype=np.float32), tmpl, opts, device=device, seeds=np.arange(S, dtype=np.int64) + 1)
    with Population(S, rad, g.chrom, g.copy_off, g.copy_beads, device=device) as pop:
        pop.upload(crd)
        hic = {"pairs": _freq_pairs(pop, g, n_pairs), "contact_range": 2.0, "k": 1.0}
        dip = np.nonzero(np.diff(g.copy_off) == 2)[0]
        probes = rng.choice(dip, 300, replace=False).astype(np.int32)
        pairs = rng.choice(dip, 1000, replace=False).astype(np.int32).reshape(500, 2)
        rlo, rhi, _, _ = pop.fish_ranks(probes)
        plo, phi, _, _ = pop.fish_ranks(pairs[:, 0], pairs[:, 1])
    noisy = lambda v: np.sort(np.abs(v.astype(np.float64) + rng.normal(0, 50.0, v.shape)) + 100.0, axis=1)
    fish = {"rtype": "rRpP", "tol": 25.0, "k": 1.0, "center": n + 2, "probes": probes, "pairs": pairs,
            "radial_min": noisy(rlo), "radial_max": noisy(rhi), "pair_min": noisy(plo), "pair_max": noisy(phi)}
    prof = np.clip(rng.normal(0.3, 0.25, g.n_hap), 0, 1)
    loci = np.nonzero(prof >= 0.3)[0].astype(np.int32)                       # the last sigma of igm-config_all.json
    damid = {"loci": loci, "pwish": prof[loci], "plast": np.zeros(loci.size), "contact_range": 0.05, "shape": "ellipsoid",
             "nucleus_param": SEMI, "k": 1.0}
    t0 = time.perf_counter()
    out = iteration.am_iteration(crd, g, rad, tmpl, opts, hic=hic, damid=damid, fish=fish, device=device,
                                 seeds=np.arange(S, dtype=np.int64) + 7, violations=True)
    dt = time.perf_counter() - t0
    asm = out["restraints"]
    v = np.array(out["violations"], dtype=np.int64)
    return {"config": "configs[2]: 2 Mb diploid Hi-C + lamina DamID + ellipsoidal envelope + FISH pairs/radial (rRpP)",
            "structures": S, "beads": n, "seconds": dt, "structures_per_s": S / dt, "astep_seconds": out["timings"]["astep_s"],
            "mstep_seconds": out["timings"]["mstep_s"], "md_kernel_ms": out["timings"]["md_kernel_ms"],
            "hic_pairs": int(hic["pairs"][0].size), "hic_bonds_per_structure": float(np.bitwise_count(asm["hic_bits"]).sum()) / S,
            "damid_loci": int(loci.size), "damid_members_per_structure": float(np.bitwise_count(asm["damid_bits"]).sum()) / S,
            "fish_bonds_per_structure": int(asm["fish_bonds"][0][1]), "restraints_imposed": int(v[:, 0].sum() + v[:, 2].sum()),
            "restraints_violated": int(v[:, 1].sum() + v[:, 3].sum())}


def run_c4(device=0, S=8, n_clusters=2000):
    g = synth.genome_200kb()
    n = len(g)
    rad = synth.bead_radii(n, 5500.0, 0.4)
    opts = synth.demo_protocol_options()
    rng = np.random.default_rng(13)
    start = synth.population_territories(g, S, R=5500.0, seed=777)
    tmpl = synth.pack_structure(np.zeros((n, 3), dtype=np.float32), rad, g, nucleus_radius=5500.0, seed=0, struct_id=0)
    t0 = time.perf_counter()
    crd, _ = iteration.relax_init(start, tmpl, opts, device=device, seeds=np.arange(S, dtype=np.int64) + 1)
    t_relax = time.perf_counter() - t0
    # Hi-C: power law within a chromosome, neighbours up to 40 bins apart
    chrom = g.chrom[:g.n_hap]
    i = np.repeat(np.arange(g.n_hap), 40)
    j = i + np.tile(np.arange(1, 41), g.n_hap)
    ok = (j < g.n_hap)
    i, j = i[ok], j[ok]
    ok = chrom[i] == chrom[j]
    i, j = i[ok], j[ok]
    pw = np.minimum(1.0, 0.9 * (j - i).astype(np.float64) ** -1.08).astype(np.float32)
    keep = pw >= np.float32(0.05)
    pairs = (i[keep].astype(np.int32), j[keep].astype(np.int32), pw[keep].astype(np.float64), np.zeros(int(keep.sum())))
    # SPRITE clusters: sizes 2 + Geom(0.3), loci within +-5 Mb windows (25 bins), at most 2 chromosomes
    clusters = []
    for _ in range(n_clusters):
        size = 2 + rng.geometric(0.3)
        c0 = rng.integers(0, g.n_hap)
        win = np.arange(max(0, c0 - 25), min(g.n_hap, c0 + 26))
        clusters.append(rng.choice(win, min(size, win.size), replace=False).astype(np.int64))
    rs = np.random.RandomState(13)
    t0 = time.perf_counter()
    batch = sprite.ClusterBatch(clusters, chrom, g.copy_off, g.copy_beads, rs, max_chrom=6)
    with Population(S, rad, g.chrom, g.copy_off, g.copy_beads, device=device) as pop:
        pop.upload(crd)
        t1 = time.perf_counter()
        best_rg2, best_struct, selected = pop.sprite_select(batch, keep_best=min(50, S - 1))
        sprite_ms = pop.last_kernel_ms()
        rows = pop.actdist(*pairs, 2.0)
        actdist_ms = pop.last_kernel_ms()
        bits = pop.assign(rows[0], rows[1], rows[2])
    assignment, pos = sprite.assign_clusters(best_rg2, best_struct, S, rs, batch_size=10, kT=50.0)
    beads = [selected[c][pos[c]] for c in range(len(clusters))]
    atoms, bonds = sprite.centroid_restraints(crd, rad, assignment, beads, 0, tmpl.n_atoms, 0.005, k=1.0)
    t_astep = time.perf_counter() - t0
    extras = rp.extras(S, bonds=bonds, atoms=atoms)
    r0 = (2.0 * (rad[rows[0]] + rad[rows[1]])).astype(np.float64)
    prot = _abi.make_protocol(opts)
    t0 = time.perf_counter()
    new, results, _, tim = run_population(tmpl, crd, n, rows[0], rows[1], r0, 1.0, bits, prot, device=device,
                                          seeds=np.arange(S, dtype=np.int64) + 7, extras=extras)
    t_mstep = time.perf_counter() - t0
    evals = sum(prot.mdsteps[k] + 1 + prot.relax_mdsteps + 1 for k in range(prot.n_stage))
    assert all(r.status == 0 for r in results)
    return {"config": "configs[3]: 200 kb diploid Hi-C + SPRITE through the cluster-per-structure MD kernel",
            "structures": S, "beads": n, "seconds": t_astep + t_mstep, "astep_seconds": t_astep, "mstep_seconds": t_mstep,
            "md_kernel_ms": tim["md_kernel_ms"], "cg_kernel_ms": tim["cg_kernel_ms"], "md_force_evals": evals,
            "us_per_force_eval_and_structure": tim["md_kernel_ms"] * 1e3 / evals,
            "cluster_ctas": 16 if S <= 7 else 8,      # upload()'s choice at this size: a B200 holds 7 clusters of 16 CTAs or 15 of 8 at once
            "hic_pairs": int(pairs[0].size), "hic_bonds_per_structure": float(np.bitwise_count(bits).sum()) / S,
            "sprite_clusters": n_clusters, "sprite_select_kernel_ms": sprite_ms, "actdist_kernel_ms": actdist_ms,
            "sprite_centroids_per_structure": float(atoms[0][-1]) / S, "relax_init_seconds": t_relax,
            "structures_per_s": S / t_mstep}


if __name__ == "__main__":
    which = sys.argv[1:] or ["c3", "c4"]
    for w in which:
        print(json.dumps({"c3": run_c3, "c4": run_c4}[w]()), flush=True)
