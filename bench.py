#!/usr/bin/env python
"""Benchmark of the A/M hot path on B200 (contract: see DESIGN.md "Measurement").

    python bench.py --gpus N --steps K --warmup W [--impl reference]

Workload (BASELINE.json configs[1]): 2 Mb diploid genome (3008 beads + the envelope's dummy),
1000 structures PER GPU (weak scaling), spherical envelope, Hi-C restraints assigned at one
sigma, the demo's full annealing protocol (4 x [relax 500 + 5k/15k/15k/10k MD steps] + CG).
A "step" = one M-step over the rank's structures.  `value` = structures/s with everything
resident in HBM; `e2e` = the same from HOST buffers through igm_b200.iteration.mstep_population:
coordinates + actdist rows up, contact assignment and the assembly of every structure on the device
(igmk_submit_population_dev), MD + CG, optimised coordinates back; `e2e_host_assembly` = the round-1 form,
igm_b200.kernel.run_population with the assignment bit matrix on the host (chunked submit -> upload -> run ->
fetch over the C ABI, chunks pipelined on streams); `e2e_plugin` = per-structure optimize(model, cfg) calls
under B200Controller.  The A-step (activation distances + assignment over all structures x pairs) is
timed in the same run and reported under "astep"; "am_iteration" is a whole A/M iteration from host coordinates.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from igm_b200 import _abi, synth, parallel          # noqa: E402

NUCLEUS_R = 5500.0
FP32_PEAK_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12      # non-tensor FP32 FMA peak at max clock (derived; fallback when the probe fails)
FIXTURE = os.path.join(ROOT, "tests", "golden", "bench_c2_structures.npz")
N_FIXTURE = 32                 # structures of the default workload kept as a fixture: the reference arm's sample


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--structures", type=int, default=1000, help="structures per GPU")
    ap.add_argument("--pairs", type=int, default=614140,
                    help="haploid pairs kept by the sigma threshold (default: what sigma=0.01 keeps of the demo matrix)")
    ap.add_argument("--md-scale", type=float, default=1.0, help="scale every mdsteps entry (diagnostics only)")
    ap.add_argument("--relax-md-scale", type=float, default=None, help="mdsteps scale of the untimed RelaxInit pass (profiling aid)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-sample", type=int, default=0, help="structures in the CPU baseline sample (0: one per core)")
    ap.add_argument("--write-fixture", default=None,
                    help="write the first %d structures of the workload (M-step inputs: coordinates + assigned Hi-C bonds) to this "
                         ".npz; commit it as tests/golden/bench_c2_structures.npz" % N_FIXTURE)
    ap.add_argument("--no-plugin", action="store_true", help="skip the e2e_plugin leg (B200Controller + optimize(model, cfg))")
    ap.add_argument("--no-extra", action="store_true", help="skip the configs[2] / configs[3] iterations (tools/bench_configs.py)")
    return ap.parse_args()


def workload_config(args, evals):
    """The `config` object of the JSON line.  Both arms print exactly this (the driver compares them): it holds only what
    DEFINES the workload; what a run measures about it (sigma, bonds per structure, ...) goes to `workload_stats`."""
    return {"workload": "configs[1]: 2 Mb diploid Hi-C, %d structures per GPU, spherical envelope (R = 5500), full demo annealing "
                        "protocol + CG; Hi-C matrix = contact frequencies of the relaxed synthetic population of rank 0, "
                        "thresholded to keep %d haploid pairs (the demo matrix keeps 614140 at its last sigma, 0.01)"
                        % (args.structures, args.pairs),
            "beads": 3008, "structures_per_gpu": args.structures, "hic_pairs": args.pairs, "md_force_evals": evals,
            "md_scale": args.md_scale,
            "fixture": "tests/golden/bench_c2_structures.npz = M-step inputs (relaxed coordinates + assigned Hi-C bonds) of "
                       "structures 0..%d of this workload; the reference arm and cpu_baseline run those" % (N_FIXTURE - 1),
            "l2_policy": "working set per step (bond lists + neighbour tables, >400 MB) exceeds L2; the A-step slab is L2 "
                         "resident at 1000 structures (36 MB): roofline_astep says so"}


def load_fixture():
    if not os.path.isfile(FIXTURE):
        return None
    z = np.load(FIXTURE)
    return {k: z[k] for k in z.files}


def fixture_systems(fx, g, rad, n):
    """PackedSystems of the first n fixture structures, assembled like the GPU arm's (same seeds, same bond order)."""
    out = []
    for s in range(min(n, fx["crd"].shape[1])):
        lo, hi = int(fx["bond_off"][s]), int(fx["bond_off"][s + 1])
        bi, bj = fx["bond_i"][lo:hi].astype(np.int32), fx["bond_j"][lo:hi].astype(np.int32)
        r0 = (2.0 * (rad[bi] + rad[bj])).astype(np.float64)
        out.append(synth.pack_structure(fx["crd"][:, s, :], rad, g, bi, bj, r0, nucleus_radius=NUCLEUS_R, seed=7 + s, struct_id=s))
    return out


# --------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        super(ClockSampler, self).__init__(daemon=True)
        self.device, self.rows, self._halt = device, [], threading.Event()

    def run(self):
        while not self._halt.is_set():
            try:
                out = subprocess.check_output(["nvidia-smi", "-i", str(self.device), "--query-gpu=" + self.Q,
                                               "--format=csv,noheader,nounits"], timeout=10).decode()
                self.rows.append([v.strip() for v in out.strip().split(",")])
            except Exception:
                pass
            self._halt.wait(0.5)

    def finish(self):
        self._halt.set()
        self.join(timeout=15)
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(r[2 + k].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
                "samples": len(self.rows)}


def protocol(md_scale):
    opts = synth.demo_protocol_options()
    if md_scale != 1.0:
        cap = opts["custom_annealing_protocol"]
        cap["mdsteps"] = [max(1, int(v * md_scale)) for v in cap["mdsteps"]]
        cap["relax"]["mdsteps"] = max(1, int(cap["relax"]["mdsteps"] * md_scale))
    return opts, _abi.make_protocol(opts)


def md_force_evals(prot):
    """force evaluations of the MD launch per structure: every run = setup + nsteps"""
    n = 0
    for k in range(prot.n_stage):
        if prot.relax_mdsteps > 0:
            n += prot.relax_mdsteps + 1
        n += prot.mdsteps[k] + 1
    return n


SKIN = 0.7            # neighbour skin of k_md2 in bead radii (igmk.cu: skin_factor)
SMEM_PEAK_TBS = 148 * 128 * 1.965e9 / 1e12          # 128 B per clock and SM at max clock (derived)


def work_per_eval(systems, coords, sample=6):
    """Work of one force evaluation of one structure, counted on the optimised coordinates of a few structures.
    -> (algorithmic flops, executed flops, executed shared-memory bytes)
    algorithmic: SURVEY.md 8(d), F_eval = 8 P_cand + 21 P_act + 20 B + 9 N_env + 36 N with every pair and every
      bond counted ONCE (half list; P_cand = pairs inside cutoff + the kernel's skin);
    executed: k_md2 gathers, so it walks every pair and every bond from both ends (2 P_cand, 2 P_act, 2 B);
    shared memory: one 16-byte position read per neighbour-table entry and per bond item, plus the atom's own."""
    from scipy.spatial import cKDTree
    alg, exe, smem = [], [], []
    for ps, xyz in list(zip(systems, coords))[:sample]:
        n = ps.n_atoms - 1
        r = float(ps.radius[0])
        t = cKDTree(xyz[:n].astype(np.float64))
        p_cand = (t.count_neighbors(t, (2.0 + SKIN) * r) - n) // 2     # unordered pairs within cutoff + skin
        p_act = (t.count_neighbors(t, 2.0 * r) - n) // 2
        b = ps.bond_i.size
        alg.append(8.0 * p_cand + 21.0 * p_act + 20.0 * b + 9.0 * n + 36.0 * n)
        exe.append(2 * (8.0 * p_cand + 21.0 * p_act + 20.0 * b) + 9.0 * n + 36.0 * n)
        smem.append(16.0 * (2 * p_cand + 2 * b + n))
    return float(np.mean(alg)), float(np.mean(exe)), float(np.mean(smem))


# --------------------------------------------------------------------------------------
def build_workload(args, rank, world, device, log):
    """-> dict(genome, radii, systems (local PackedSystems), crd_local, pairs, rows, timings of the A-step)"""
    import torch
    import torch.distributed as dist
    from igm_b200.kernel import MStepBatch
    from igm_b200.astep import Population

    S = args.structures
    g = synth.genome_2mb()
    n = len(g)
    rad = synth.bead_radii(n, NUCLEUS_R, 0.4)
    opts, prot = protocol(args.md_scale)
    t0 = time.time()
    start = np.empty((n, S, 3), dtype=np.float32)
    for s in range(S):
        start[:, s, :] = synth.territories(g, NUCLEUS_R, np.random.default_rng(1234 + rank * S + s))
    # RelaxInit-equivalent pass (igm/steps/RelaxInit.py:86-177: polymer + steric + envelope only)
    relax_sys = [synth.pack_structure(start[:, s, :], rad, g, nucleus_radius=NUCLEUS_R, seed=1 + rank * S + s,
                                      struct_id=rank * S + s) for s in range(S)]
    crd = np.empty((n, S, 3), dtype=np.float32)
    relax_prot = prot if args.relax_md_scale is None else protocol(args.relax_md_scale)[1]
    with MStepBatch(device) as b:
        slots = [b.submit(ps) for ps in relax_sys]
        b.run(relax_prot)
        for s, slot in enumerate(slots):
            crd[:, s, :] = b.fetch(slot)[0][:n]
        relax_ms = sum(b.last_timings()[:2])
    log("setup: territories + RelaxInit pass of %d structures: %.1f s (kernels %.0f ms)" % (S, time.time() - t0, relax_ms))

    # ---- A-step: all-gather the slab on the struct axis, shard the pairs ----
    local = torch.from_numpy(crd).to("cuda:%d" % device)
    slab = parallel.all_gather_slab(local)            # NCCL all-gather on the structure axis
    S_tot = S * world
    astep = {}
    # Synthetic Hi-C matrix that is geometrically consistent with the population (a random matrix
    # collapses the genome into a globule): the contact frequency of every haploid pair in the relaxed
    # structures of RANK 0 is the "experiment" (the same matrix, sigma and pair list whatever the number of
    # GPUs, so the restraint load per structure does not drift with N); the iteration wishes 10 % more
    # contacts than are present (pwish = 1.1 f) and the previous iteration is taken to have imposed f (plast).
    iu, ju = np.triu_indices(g.n_hap, 1)
    rank0 = slab[:, :S, :].contiguous()
    with Population(S, rad, g.chrom, g.copy_off, g.copy_beads, device=device) as pop0:
        pop0.attach(rank0)
        freq = pop0.contact_freq(iu.astype(np.int32), ju.astype(np.int32), 2.0)
    del rank0
    with Population(S_tot, rad, g.chrom, g.copy_off, g.copy_beads, device=device) as pop:
        pop.attach(slab)
        pw_all = np.minimum(np.float32(1.0), np.float32(1.1) * freq).astype(np.float32)
        sigma = float(np.partition(pw_all, pw_all.size - args.pairs)[pw_all.size - args.pairs]) if args.pairs < pw_all.size else 0.0
        keep = pw_all >= np.float32(sigma)
        log("synthetic Hi-C matrix: pairs with pwish >= sigma: " + ", ".join(
            "%g: %d" % (sg, int((pw_all >= np.float32(sg)).sum())) for sg in (0.05, 0.02, 0.01, 0.005, 0.003, 0.002, 0.001)))
        pi, pj = iu[keep].astype(np.int32), ju[keep].astype(np.int32)
        pw = pw_all[keep].astype(np.float64)
        pl = np.round(freq[keep].astype(np.float64), 4).astype(np.float32).astype(np.float64)
        lo_p, hi_p = parallel.pair_shards(parallel.pair_cost(pi, pj, g.chrom, g.copy_off), world)[rank]
        sl = slice(lo_p, hi_p)
        kt, et = [], []
        rows = None
        for rep in range(3):
            torch.cuda.synchronize(device)
            t1 = time.perf_counter()
            rows = pop.actdist(pi[sl], pj[sl], pw[sl], pl[sl], 2.0)     # this rank's contiguous chunk of the pair list
            et.append(time.perf_counter() - t1)
            kt.append(pop.last_kernel_ms())
        astep["actdist_kernel_ms"] = min(kt)
        astep["actdist_e2e_ms"] = min(et) * 1e3
        astep["pairs_local"] = int(pi[sl].size)
        n_copies = np.diff(g.copy_off)
        astep["actdist_bytes"] = float(12.0 * S_tot * (n_copies[pi[sl]] + n_copies[pj[sl]]).sum() + 32.0 * pi[sl].size)
    rows = parallel.all_gather_rows(rows, device="cuda:%d" % device)      # rank order == global pair order
    row, col, dst, prb = rows
    # the structure-sharded variant (configs[4]: nobody holds the whole slab; per-pair counts and radix histograms are
    # all-reduced over NCCL): timed on the same pairs, and its rows must equal the gathered ones bit for bit
    with Population(S, rad, g.chrom, g.copy_off, g.copy_beads, device=device) as pop:
        pop.attach(local)
        torch.cuda.synchronize(device)
        t1 = time.perf_counter()
        rows_sh = parallel.sharded_actdist(pop, S_tot, pi, pj, pw, pl, 2.0)
        astep["sharded_e2e_ms"] = (time.perf_counter() - t1) * 1e3
        astep["sharded_kernel_ms"] = parallel.sharded_actdist.last_kernel_ms
        astep["sharded_identical"] = bool(all(np.array_equal(a.view(np.uint32) if a.dtype == np.float32 else a,
                                                             b.view(np.uint32) if b.dtype == np.float32 else b)
                                              for a, b in zip(rows_sh, rows)))
    del rows_sh
    with Population(S, rad, g.chrom, g.copy_off, g.copy_beads, device=device) as pop:
        pop.attach(local)
        kt, et = [], []
        for rep in range(3):
            t1 = time.perf_counter()
            bits = pop.assign(row, col, dst)
            et.append(time.perf_counter() - t1)
            kt.append(pop.last_kernel_ms())
        astep["assign_kernel_ms"] = min(kt)
        astep["assign_e2e_ms"] = min(et) * 1e3
        astep["rows"] = int(row.size)
        astep["assign_bytes"] = float(row.size) * (24.0 * S + 12.0 + 4.0 * ((S + 31) // 32))
    # bits -> per-structure Hi-C bonds (HarmonicUpperBound(i, j, contact_range*(ri+rj), k), inter_hic.py:55-62):
    # done for the whole population by igmk_submit_population; a few structures are also packed here in
    # Python for the CPU baseline and the flop count
    r0_all = (2.0 * (rad[row] + rad[col])).astype(np.float64)
    n_py = min(S, max(os.cpu_count() or 1, N_FIXTURE))
    systems, masks = [], []
    for s in range(n_py):
        m = ((bits[:, s >> 5] >> np.uint32(s & 31)) & np.uint32(1)).astype(bool)
        masks.append(m)
        systems.append(synth.pack_structure(crd[:, s, :], rad, g, row[m], col[m], r0_all[m], nucleus_radius=NUCLEUS_R,
                                            seed=7 + rank * S + s, struct_id=rank * S + s))
    # ---- the fixture the reference arm runs: M-step inputs of this rank-0 workload's first structures ----
    fixture_check = None
    if rank == 0:
        nf = min(N_FIXTURE, n_py)
        if args.write_fixture:
            off = np.concatenate([[0], np.cumsum([int(m.sum()) for m in masks[:nf]])]).astype(np.int64)
            np.savez_compressed(args.write_fixture, crd=crd[:, :nf, :].copy(), bond_off=off,
                                bond_i=np.concatenate([row[m] for m in masks[:nf]]).astype(np.uint16),
                                bond_j=np.concatenate([col[m] for m in masks[:nf]]).astype(np.uint16),
                                structures=np.int64(S), pairs=np.int64(args.pairs), sigma=np.float64(sigma))
            log("fixture written: %s (%d structures)" % (args.write_fixture, nf))
        fx = load_fixture()
        if fx is None:
            fixture_check = {"present": False}
        else:
            nf = min(nf, fx["crd"].shape[1])
            same_crd = int(sum(np.array_equal(fx["crd"][:, s, :], crd[:, s, :]) for s in range(nf)))
            same_bonds, fx_counts = 0, []
            for s in range(nf):
                lo, hi = int(fx["bond_off"][s]), int(fx["bond_off"][s + 1])
                fx_counts.append(hi - lo)
                same_bonds += int(np.array_equal(fx["bond_i"][lo:hi], row[masks[s]].astype(np.uint16)) and
                                  np.array_equal(fx["bond_j"][lo:hi], col[masks[s]].astype(np.uint16)))
            mine = [int(m.sum()) for m in masks[:nf]]
            fixture_check = {"present": True, "structures": nf, "coordinates_bit_identical": same_crd, "bond_lists_identical": same_bonds,
                             "hic_bonds_fixture_mean": float(np.mean(fx_counts)), "hic_bonds_this_run_mean": float(np.mean(mine)),
                             "comparable": bool(int(fx["structures"]) == S and int(fx["pairs"]) == args.pairs and world >= 1)}
            log("fixture check: %s" % json.dumps(fixture_check))
    nbonds = float(np.bitwise_count(bits).sum()) / S if hasattr(np, "bitwise_count") else float(np.unpackbits(bits.view(np.uint8)).sum()) / S
    log("A-step: %d pairs -> %d rows; actdist %.2f ms, assign %.2f ms (kernels); %.0f Hi-C bonds/structure"
        % (pi.size, row.size, astep["actdist_kernel_ms"], astep["assign_kernel_ms"], nbonds))
    cnts = np.sort(np.bincount(np.concatenate([systems[0].bond_i, systems[0].bond_j]), minlength=3072))[::-1]
    cnts = np.concatenate([cnts, np.zeros((-cnts.size) % 32, dtype=cnts.dtype)]).reshape(-1, 32)
    log("bond items of structure 0: %d real, %d after sliced-ELL padding, longest row %d"
        % (cnts.sum(), 32 * cnts.max(axis=1).sum(), cnts.max()))
    astep["pairs_total"] = int(pi.size)
    astep["sigma"] = sigma
    astep["hic_bonds_per_structure"] = nbonds
    template = synth.pack_structure(np.zeros((n, 3), dtype=np.float32), rad, g, nucleus_radius=NUCLEUS_R, seed=0, struct_id=rank * S)
    seeds = (7 + rank * S + np.arange(S)).astype(np.int64)
    return {"genome": g, "radii": rad, "systems": systems, "crd": crd, "astep": astep, "prot": prot, "opts": opts,
            "template": template, "seeds": seeds, "rows": (row, col, dst, r0_all), "bits": bits, "pairs": (pi, pj, pw, pl),
            "sigma": sigma, "fixture_check": fixture_check, "slab_structures": S_tot}


def cpu_mstep_baseline(systems, prot, n_sample, threads):
    """the oracle's C port of the whole protocol, one structure per host thread (ctypes drops the GIL)"""
    from concurrent.futures import ThreadPoolExecutor
    from oracle import ff
    ff.lib()
    sample = systems[:n_sample]
    t0 = time.perf_counter()
    with ThreadPoolExecutor(max_workers=threads) as ex:
        res = list(ex.map(lambda ps: ff.optimize(ps, prot)[1].final_energy, sample))
    dt = time.perf_counter() - t0
    return len(sample) / dt, dt, res


def cpu_astep_baseline(wl, n_pairs=2000, budget_s=12.0):
    """The A-step restatement (oracle/astep_oracle.py = get_actdist + the text round trip, the reference's numpy code
    line for line) on the host, on a bounded sample of the bench's own pair list over the same population:
    one core, then all cores (one process per core, like the reference's ipyparallel map over batches)."""
    from concurrent.futures import ProcessPoolExecutor
    from oracle import astep_oracle
    g, rad, crd = wl["genome"], wl["radii"], wl["crd"]
    pi, pj, pw, pl = wl["pairs"]
    sel = np.linspace(0, pi.size - 1, min(n_pairs, pi.size)).astype(np.int64)      # spread over the list: intra and inter pairs
    a = (pi[sel], pj[sel], pw[sel], pl[sel])
    t0 = time.perf_counter()
    n1 = 0
    for lo in range(0, sel.size, 100):                                              # single core, until the budget is spent
        astep_oracle.actdist_rows(a[0][lo:lo + 100], a[1][lo:lo + 100], a[2][lo:lo + 100], a[3][lo:lo + 100], crd, rad,
                                  g.chrom, g.copy_index, 2.0)
        n1 += min(100, sel.size - lo)
        if time.perf_counter() - t0 > budget_s / 2:
            break
    one = n1 / (time.perf_counter() - t0)
    cores = os.cpu_count() or 1
    _ASTEP_SHARED.update(crd=crd, rad=rad, chrom=g.chrom, ci=g.copy_index)
    n_all = int(max(sel.size, min(pi.size, one * cores * budget_s / 2)))            # about budget_s / 2 seconds on all cores
    sel2 = np.linspace(0, pi.size - 1, min(n_all, pi.size)).astype(np.int64)
    chunks = [sel2[k::cores] for k in range(cores)]
    import multiprocessing as mp
    t0 = time.perf_counter()
    with ProcessPoolExecutor(max_workers=cores, mp_context=mp.get_context("fork")) as ex:
        list(ex.map(_astep_chunk, [(pi[c], pj[c], pw[c], pl[c]) for c in chunks]))
    dt = time.perf_counter() - t0
    return {"value": sel2.size / dt, "unit": "pairs/s", "cores": cores, "kind": "port", "one_core": one,
            "sample": "%d of the %d haploid pairs over all %d structures on %d processes (%.1f s); one core: %d pairs (%.0f pairs/s)"
                      % (sel2.size, pi.size, crd.shape[1], cores, dt, n1, one)}


_ASTEP_SHARED = {}


def _astep_chunk(a):
    from oracle import astep_oracle
    d = _ASTEP_SHARED
    return astep_oracle.actdist_rows(a[0], a[1], a[2], a[3], d["crd"], d["rad"], d["chrom"], d["ci"], 2.0)[0].size


def run_reference(args, rank, world):
    """--impl reference: the CPU restatement of the reference path (lmp_serial itself is not installable here, see
    DESIGN.md) on all host cores, on the GPU arm's own workload: every step optimises the fixture structures -- the
    M-step inputs (relaxed coordinates + assigned Hi-C bonds) of the first structures of the default workload, written
    by `bench.py --write-fixture` on a B200 -- one structure per host thread, full protocol."""
    if rank != 0:
        return
    g = synth.genome_2mb()
    n = len(g)
    rad = synth.bead_radii(n, NUCLEUS_R, 0.4)
    opts, prot = protocol(args.md_scale)
    cores = os.cpu_count() or 1
    n_sample = min(args.cpu_sample or cores, N_FIXTURE)
    fx = load_fixture()
    if fx is None:
        print(json.dumps({"impl": "reference", "unavailable": "tests/golden/bench_c2_structures.npz missing: run "
                                                              "`bench.py --write-fixture` on a B200 first"}))
        return
    systems = fixture_systems(fx, g, rad, n_sample)
    n_sample = len(systems)
    n_hic = float(np.mean([ps.bond_i.size for ps in systems])) - float(synth.polymer_bonds(g, rad)[0].size)
    times = []
    for it in range(args.warmup + args.steps):
        v, dt, _ = cpu_mstep_baseline(systems, prot, n_sample, cores)
        if it >= args.warmup:
            times.append(dt)
    ms = 1e3 * float(np.mean(times))
    value = n_sample / (ms / 1e3)
    sample = ("structures 0..%d of the workload (tests/golden/bench_c2_structures.npz: relaxed coordinates + %.0f assigned Hi-C "
              "bonds per structure), one per host thread, full protocol; C restatement of the lammpgen script "
              "(oracle/ff_oracle.c) -- lmp_serial is not installable offline" % (n_sample - 1, n_hic))
    print(json.dumps({
        "impl": "reference", "metric": "optimised structures/sec at 2Mb diploid", "value": value, "unit": "structures/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, md_force_evals(prot)),
        "workload_stats": {"structures_per_step": n_sample, "hic_bonds_per_structure": n_hic},
        "cpu_baseline": {"value": value, "unit": "structures/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "structures/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


class _P(object):
    """duck-typed igm.model.Particle / Force / Model (the attributes pack_model and the reference's LammpsModel read,
    lammps_model.py:253-301)"""
    __slots__ = ("pos", "r", "ptype")

    def __init__(self, pos, r, ptype):
        self.pos, self.r, self.ptype = pos, r, ptype


class _F(object):
    __slots__ = ("ftype", "i", "j", "d", "k", "shape", "semiaxes", "particle_ids")

    def __init__(self, ftype, **kw):
        self.ftype = ftype
        for a, v in kw.items():
            setattr(self, a, v)


class _Model(object):
    def __init__(self, uid):
        self.id, self.particles, self.forces = uid, [], []


def plugin_models(wl, n_models, rank, S):
    """n_models structures of the workload as Model objects carrying the restraints ModelingStep.task applies
    (igm/steps/ModelingStep.py:184-264): Steric, Envelope (with its dummy), Polymer, assigned Hi-C bonds."""
    from igm_b200 import pack
    g, rad, crd = wl["genome"], wl["radii"], wl["crd"]
    row, col, dst, r0_all = wl["rows"]
    bits = wl["bits"]
    pb_i, pb_j, pb_r0 = synth.polymer_bonds(g, rad)
    n = len(g)
    models = []
    for s in range(n_models):
        m = _Model(rank * S + s)
        m.particles = [_P(crd[k, s, :].copy(), rad[k], pack.NORMAL) for k in range(n)]
        m.particles.append(_P(np.zeros(3, dtype=np.float32), 0.0, pack.DUMMY_STATIC))
        m.forces.append(_F(pack.EXCLUDED_VOLUME, k=1.0, particle_ids=list(range(n))))
        m.forces.append(_F(pack.ENVELOPE, shape="sphere", semiaxes=np.array([NUCLEUS_R] * 3), k=1.0, particle_ids=list(range(n))))
        m.forces += [_F(pack.HARMONIC_UPPER_BOUND, i=int(a), j=int(b), d=float(d), k=1.0) for a, b, d in zip(pb_i, pb_j, pb_r0)]
        sel = ((bits[:, s >> 5] >> np.uint32(s & 31)) & np.uint32(1)).astype(bool)
        m.forces += [_F(pack.HARMONIC_UPPER_BOUND, i=int(a), j=int(b), d=float(d), k=1.0)
                     for a, b, d in zip(row[sel], col[sel], r0_all[sel])]
        models.append(m)
    return models


def plugin_e2e(wl, n_models, rank, S, device):
    """The path IGM itself takes: controller.map(task) with one task per structure, every task calling
    optimize(model, cfg) (ModelingStep.py:97, :335) -- B200Controller parks the calls and launches them together."""
    from igm_b200 import kernel
    from igm_b200.controller import B200Controller
    models = plugin_models(wl, n_models, rank, S)
    cfg = {"optimization": {"kernel": "b200", "optimizer_options": dict(wl["opts"], seed=7), "kernel_opts": {"b200": {"device": device}}},
           "runtime": {"step_no": 1}}
    ctl = B200Controller(max_threads=n_models)
    t0 = time.perf_counter()
    infos = ctl.map(lambda k: kernel.optimize(models[k], cfg), range(n_models))
    dt = time.perf_counter() - t0
    if not all(np.isfinite(i["final-energy"]) for i in infos):
        raise RuntimeError("bench: plugin path returned a non-finite energy")
    return dt, ctl.last_batcher.batches


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    import torch
    import torch.distributed as dist
    from igm_b200.kernel import MStepBatch

    def log(msg):
        if rank == 0:
            print("[bench] " + msg, file=sys.stderr, flush=True)

    device = local_rank
    torch.cuda.set_device(device)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda:%d" % device))
    # measured FP32 FMA / shared-memory rates of this GPU (probe.cuh): the denominators of the MD kernel's roofline
    import ctypes as C
    pk = [C.c_double(0), C.c_double(0), C.c_double(0)]
    _abi.check(_abi.load().igmk_peak_probe(device, C.byref(pk[0]), C.byref(pk[1]), C.byref(pk[2])), "igmk_peak_probe")
    fp32_peak, fp32x2_peak, smem_peak = pk[0].value, pk[1].value, pk[2].value
    log("probe: FFMA %.1f TFLOP/s, FFMA2 %.1f TFLOP/s, LDS.128 %.1f TB/s" % (fp32_peak, fp32x2_peak, smem_peak))
    wl = build_workload(args, rank, world, device, log)
    systems, prot = wl["systems"], wl["prot"]
    S = args.structures
    g, rad, crd = wl["genome"], wl["radii"], wl["crd"]
    row, col, dst, r0_all = wl["rows"]

    def submit_all(b):
        b.submit_population(wl["template"], crd, len(g), row, col, r0_all, 1.0, wl["bits"], seeds=wl["seeds"])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(device)

    # ---- value: batch resident in HBM, K steps inside ONE timed region --------------------
    batch = MStepBatch(device)
    submit_all(batch)
    batch.upload()
    md_ms, cg_ms = [], []
    for _ in range(args.warmup):
        batch.reset_coordinates()
        batch.run(prot)
    sampler = ClockSampler(device)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        batch.reset_coordinates()
        batch.run(prot)
        t = batch.last_timings()
        md_ms.append(t[0]); cg_ms.append(t[1])
    e1.record()
    barrier()
    clocks = sampler.finish()
    ms_step = e0.elapsed_time(e1) / args.steps
    coords = [batch.fetch(k)[0] for k in range(min(S, 6))]
    cg_evals = float(np.mean([batch.fetch(k)[1].cg_evals for k in range(S)]))
    # ---- e2e: host buffers through the public call every step (igm_b200.kernel.run_population:
    #      chunked igmk_submit_population -> upload -> run -> fetch, the chunks pipelined on streams) ----
    from igm_b200.kernel import run_population
    d2h = S * (len(g) + 1) * 16 + S * 136
    batch.close()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out = run_population(wl["template"], crd, len(g), row, col, r0_all, 1.0, wl["bits"], prot, device=device, seeds=wl["seeds"])
        h2d_host = out[3]["h2d_bytes"]
    torch.cuda.synchronize(device)
    e2e_host_ms = (time.perf_counter() - t0) * 1e3 / args.steps
    if not all(r.status == 0 for r in out[1]):
        raise RuntimeError("bench: a structure of the end-to-end pass came back non-finite")
    host_crd = out[0]
    del out
    # ---- e2e, hand-off on the device (igm_b200.iteration.mstep_population): coordinates and actdist rows from host
    #      buffers every step, assignment + assembly + optimisation on the GPU, coordinates back ----
    from igm_b200.iteration import mstep_population
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out = mstep_population(crd, g, rad, wl["template"], row, col, dst, prot, device=device, seeds=wl["seeds"])
        h2d = out[3]["h2d_bytes"]
    torch.cuda.synchronize(device)
    e2e_ms = (time.perf_counter() - t0) * 1e3 / args.steps
    if not all(r.status == 0 for r in out[1]):
        raise RuntimeError("bench: a structure of the end-to-end pass came back non-finite")
    e2e_same = bool(np.array_equal(out[0].view(np.uint32), host_crd.view(np.uint32)))
    e2e_dev_chunks = out[3].get("assembled_on_device", 0)
    del out, host_crd
    # ---- one full A/M iteration from host coordinates: upload -> all-gather -> activation distances
    #      (this rank's pair chunk) -> gather rows -> assignment -> bond build + submit -> M-step -> fetch ----
    from igm_b200.iteration import hic_iteration
    barrier()
    it = hic_iteration(crd, g, rad, wl["template"], wl["pairs"], wl["opts"], device=device, seeds=wl["seeds"])
    torch.cuda.synchronize(device)
    iter_s, t_astep, t_build = it["timings"]["total_s"], it["timings"]["astep_s"], it["timings"]["mstep_s"]
    # ---- the north-star shape: ONE 1000-structure population over the N GPUs of the box (1000 / N per GPU) ----
    ns_S = max(1, 1000 // world)
    if world == 1 and S == 1000:
        ns_s, ns_a, ns_m = iter_s, t_astep, t_build
    else:
        sub = np.ascontiguousarray(crd[:, :min(S, ns_S), :])
        barrier()
        it2 = hic_iteration(sub, g, rad, wl["template"], wl["pairs"], wl["opts"], device=device, seeds=wl["seeds"][:sub.shape[1]])
        torch.cuda.synchronize(device)
        ns_s, ns_a, ns_m = it2["timings"]["total_s"], it2["timings"]["astep_s"], it2["timings"]["mstep_s"]
        ns_S = sub.shape[1]
    # ---- the drop-in path itself: B200Controller.map(task) -> optimize(model, cfg) per structure ----
    plug_n, plug_s, plug_batches = min(S, 256), None, None
    if not args.no_plugin:
        barrier()
        plug_s, plug_batches = plugin_e2e(wl, plug_n, rank, S, device)

    tt = torch.tensor([ms_step, e2e_ms, wl["astep"]["actdist_kernel_ms"], wl["astep"]["actdist_e2e_ms"],
                       wl["astep"]["assign_kernel_ms"], wl["astep"]["assign_e2e_ms"], iter_s, t_astep, t_build, ns_s, ns_a, ns_m,
                       plug_s or 0.0, e2e_host_ms],
                      dtype=torch.float64, device="cuda:%d" % device)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms_step, e2e_ms, ad_k, ad_e, as_k, as_e, iter_s, t_astep, t_build, ns_s, ns_a, ns_m, plug_s, e2e_host_ms = [float(v) for v in tt.tolist()]
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        traffic = {}
        for name in ("r02_traffic.json", "r01_traffic.json"):
            try:
                traffic = json.load(open(os.path.join(ROOT, "profiles", name)))
                break
            except Exception:
                pass
        hbm = peaks.get("hbm_gbs", 6650.0)
        which = "measured" if "hbm_gbs" in peaks else "fallback"
        feval, feval_exec, smem_eval = work_per_eval(systems, coords)
        evals = md_force_evals(prot)
        md_t = float(np.mean(md_ms)) * 1e-3
        md_tflops = feval * evals * S / md_t / 1e12
        a = wl["astep"]
        ad_gbs = a["actdist_bytes"] / (a["actdist_kernel_ms"] * 1e-3) / 1e9
        as_gbs = a["assign_bytes"] / (a["assign_kernel_ms"] * 1e-3) / 1e9
        slab_mb = 12.0 * len(g) * wl["slab_structures"] / 1e6
        l2_resident = slab_mb < 100.0                       # 126 MB L2: the slab is re-read from L2, not from HBM
        if not (fp32_peak > 0):
            fp32_peak, peak_src = FP32_PEAK_TFLOPS, "derived: 148 SM x 128 lanes x 2 x 1.965 GHz (probe failed)"
        else:
            peak_src = "measured in-run: dependent-chain FFMA micro-kernel, one 1024-thread CTA per SM (igmk_peak_probe)"
        mdk = [k for k in traffic if k.startswith("k_md")]
        line = {
            "metric": "optimised structures/sec at 2Mb diploid", "value": S * world / (ms_step * 1e-3), "unit": "structures/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32 (MD) + f64 (CG)", "data": "synthetic",
            "config": workload_config(args, evals),
            "workload_stats": {"sigma": a["sigma"], "pairs_kept": a["pairs_total"], "hic_bonds_per_structure": a["hic_bonds_per_structure"],
                               "cg_evals_mean": cg_evals, "fixture_check": wl["fixture_check"]},
            "e2e": {"value": S * world / (e2e_ms * 1e-3), "unit": "structures/s", "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(d2h),
                    "what": "igm_b200.iteration.mstep_population per step: coordinates + actdist rows from host buffers, contact "
                            "assignment, assembly of every structure on the device (igmk_submit_population_dev), annealing MD + CG, "
                            "coordinates back to the host",
                    "chunks_assembled_on_device": int(e2e_dev_chunks), "bit_identical_to_host_assembly": e2e_same},
            "e2e_host_assembly": {"value": S * world / (e2e_host_ms * 1e-3), "unit": "structures/s", "h2d_bytes_per_step": int(h2d_host),
                                  "d2h_bytes_per_step": int(d2h),
                                  "what": "igm_b200.kernel.run_population with the assignment bit matrix on the host: structures "
                                          "assembled by igmk_submit_population on the host threads (round-1 path)"},
            "e2e_plugin": None if args.no_plugin else {
                "value": plug_n * world / plug_s, "unit": "structures/s", "structures": plug_n, "seconds": plug_s, "launches": plug_batches,
                "what": "B200Controller.map over one task per structure, each calling igm_b200.kernel.optimize(model, cfg) on a "
                        "Model of Particle / Force objects (pack_model in Python, one launch for all parked calls, positions "
                        "written back per particle); the Model objects are built outside the timed region, as ModelingStep.task "
                        "(unchanged reference code) would"},
            "gpu_launches": 2 * args.steps,
            "am_iteration": {"seconds": iter_s, "structures_per_s": S * world / iter_s, "astep_seconds": t_astep,
                             "mstep_seconds": t_build,
                             "what": "host coordinates -> upload, all-gather, activation distances, row gather, assignment, "
                                     "assembly on the device, annealing MD + CG, coordinates back on the host"},
            "am_iteration_north_star": {"seconds": ns_s, "structures": ns_S * world, "structures_per_gpu": ns_S, "astep_seconds": ns_a,
                                        "mstep_seconds": ns_m,
                                        "what": "the same iteration with 1000 / n_gpus structures per GPU: north_star's 'full 2 Mb "
                                                "diploid 1000-structure A/M iteration' when n_gpus = 8"},
            "clocks": clocks,
            "roofline": {"bound": "fp32", "kernel": "k_md2", "achieved": md_tflops, "peak": fp32_peak, "unit": "TFLOP/s",
                         "frac": md_tflops / fp32_peak,
                         "traffic": (traffic[mdk[0]]["dram_bytes"] / traffic[mdk[0]]["structures"] / traffic[mdk[0]]["force_evals"]
                                     * S * evals) if mdk else None,
                         "traffic_note": "scaled from a short ncu --set full capture (profiles/r0*_traffic.json): bytes per structure and "
                                         "force evaluation x this launch",
                         "flops_per_force_eval": feval,
                         "flops_note": "algorithmic = SURVEY 8(d) formula, every pair and bond counted once",
                         "executed": {"flops_per_force_eval": feval_exec,
                                      "tflops": feval_exec * evals * S / md_t / 1e12,
                                      "frac": feval_exec * evals * S / md_t / 1e12 / fp32_peak},
                         "smem": {"bytes_per_force_eval": smem_eval, "achieved": smem_eval * evals * S / md_t / 1e12,
                                  "peak": smem_peak if smem_peak > 0 else SMEM_PEAK_TBS, "unit": "TB/s",
                                  "frac": smem_eval * evals * S / md_t / 1e12 / (smem_peak if smem_peak > 0 else SMEM_PEAK_TBS),
                                  "peak_source": "measured in-run: conflict-free LDS.128 micro-kernel" if smem_peak > 0 else "derived",
                                  "note": "16-byte position reads the force loops issue, before bank-conflict replays"},
                         "us_per_force_eval": md_t * 1e6 / evals / max(1.0, S / 148.0),
                         "kernel_ms": md_t * 1e3, "cg_kernel_ms": float(np.mean(cg_ms)),
                         "peak_source": peak_src, "fp32x2_peak": fp32x2_peak,
                         "derived_peak": FP32_PEAK_TFLOPS},
            "astep": {"metric": "A-step haploid pairs/sec", "pairs": a["pairs_total"], "rows": a["rows"],
                      "structures": S * world,
                      "value": a["pairs_total"] / (ad_k * 1e-3), "e2e": a["pairs_total"] / (ad_e * 1e-3), "unit": "pairs/s",
                      "actdist_kernel_ms": ad_k, "actdist_e2e_ms": ad_e, "assign_kernel_ms": as_k, "assign_e2e_ms": as_e,
                      "assign_rows_x_structures_per_s": a["rows"] * float(S) / (as_k * 1e-3),
                      "structure_sharded": {"what": "configs[4] variant: the population stays sharded by structure, per-pair contact "
                                                    "counts and four radix-256 histograms are all-reduced (NCCL) instead of gathering the slab",
                                            "e2e_ms": a["sharded_e2e_ms"], "kernel_ms_this_rank": a["sharded_kernel_ms"],
                                            "rows_bit_identical_to_gathered": a["sharded_identical"]}},
            "roofline_astep": {"bound": "hbm", "kernel": "k_actdist", "achieved": ad_gbs, "peak": hbm, "unit": "GB/s",
                               "frac": ad_gbs / hbm, "l2_resident": l2_resident, "slab_mb": slab_mb,
                               "note": ("the %.0f MB coordinate slab stays in the 126 MB L2 at this size: `achieved` is algorithmic bytes "
                                        "per second served mostly from L2, not an HBM fraction; the HBM-streaming figure is the one at "
                                        ">= 8000 structures (SCALE N = 8, tools/prof_astep.py)" % slab_mb) if l2_resident else
                                       "the slab exceeds L2: coordinate rows stream from HBM",
                               "traffic": (traffic["k_actdist_reg<4>"]["dram_bytes"] / traffic["k_actdist_reg<4>"]["pairs"] * a["pairs_local"])
                               if ("k_actdist_reg<4>" in traffic and S * world == 1000) else None,
                               "traffic_note": "ncu --set full capture at 1000 structures scaled by pairs",
                               "peak_source": which + " copy bandwidth",
                               "assign_achieved": as_gbs, "assign_frac": as_gbs / hbm},
        }
        if not args.no_cpu_baseline and world == 1:
            cores = os.cpu_count() or 1
            n_sample = min(args.cpu_sample or cores, len(systems))
            v, dt, _ = cpu_mstep_baseline(systems, prot, n_sample, cores)
            line["cpu_baseline"] = {"value": v, "unit": "structures/s", "cores": cores, "kind": "port",
                                    "sample": "structures 0..%d of the %d (same restraints, full protocol), one per host "
                                              "thread, %.1f s" % (n_sample - 1, S, dt)}
            try:
                line["astep"]["cpu_baseline"] = cpu_astep_baseline(wl)
            except Exception as e:          # the A-step baseline must never cost the bench line
                line["astep"]["cpu_baseline"] = {"error": repr(e)}
        if world == 1 and not args.no_extra:
            # BASELINE configs[2] and configs[3] at their genome sizes on a reduced number of structures (tools/bench_configs.py)
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            line["extra_configs"] = {}
            try:
                import bench_configs
                for name, fn in (("c3", bench_configs.run_c3), ("c4", bench_configs.run_c4)):
                    try:
                        line["extra_configs"][name] = fn(device)
                    except Exception as e:                      # never at the cost of the headline line
                        line["extra_configs"][name] = {"error": repr(e)}
            except Exception as e:
                line["extra_configs"] = {"error": repr(e)}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
