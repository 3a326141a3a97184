#!/usr/bin/env python
"""Benchmark of the A/M hot path on B200 (contract: see DESIGN.md "Measurement").

    python bench.py --gpus N --steps K --warmup W [--impl reference]

Workload (BASELINE.json configs[1]): 2 Mb diploid genome (3008 beads + the envelope's dummy),
1000 structures PER GPU (weak scaling), spherical envelope, Hi-C restraints assigned at one
sigma, the demo's full annealing protocol (4 x [relax 500 + 5k/15k/15k/10k MD steps] + CG).
A "step" = one M-step over the rank's structures.  `value` = structures/s with everything
resident in HBM; `e2e` = the same from host buffers through igm_b200.kernel.run_population
(chunked submit -> upload -> run -> fetch over the C ABI, chunks pipelined on streams).  The A-step (activation distances + assignment over all structures x pairs) is
timed in the same run and reported under "astep".
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
REF_HIC_BONDS = 27588          # Hi-C bonds per structure of the GPU arm's default workload (its config.hic_bonds_per_structure)
FP32_PEAK_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12      # non-tensor FP32 FMA peak at max clock (derived)


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
    return ap.parse_args()


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
    with Population(S_tot, rad, g.chrom, g.copy_off, g.copy_beads, device=device) as pop:
        pop.attach(slab)
        # Synthetic Hi-C matrix that is geometrically consistent with the population (a random matrix
        # collapses the genome into a globule): the contact frequency of every haploid pair in the relaxed
        # population is the "experiment"; the iteration wishes 10 % more contacts than are present
        # (pwish = 1.1 f) and the previous iteration is taken to have imposed f (plast).
        iu, ju = np.triu_indices(g.n_hap, 1)
        freq = pop.contact_freq(iu.astype(np.int32), ju.astype(np.int32), 2.0)
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
    n_py = min(S, max(os.cpu_count() or 1, 16))
    systems = []
    for s in range(n_py):
        m = ((bits[:, s >> 5] >> np.uint32(s & 31)) & np.uint32(1)).astype(bool)
        systems.append(synth.pack_structure(crd[:, s, :], rad, g, row[m], col[m], r0_all[m], nucleus_radius=NUCLEUS_R,
                                            seed=7 + rank * S + s, struct_id=rank * S + s))
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
            "sigma": sigma}


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


def run_reference(args, rank, world):
    """--impl reference: the CPU restatement of the reference path (lmp_serial itself is not
    installable here) on all host cores, same config / metric."""
    if rank != 0:
        return
    g = synth.genome_2mb()
    n = len(g)
    rad = synth.bead_radii(n, NUCLEUS_R, 0.4)
    opts, prot = protocol(args.md_scale)
    cores = os.cpu_count() or 1
    n_sample = args.cpu_sample or cores
    # The GPU arm derives its Hi-C restraints from a relaxed population with its own kernels; this arm may not touch
    # them, so every structure gets the same NUMBER of Hi-C upper bounds (what the GPU arm reports as
    # hic_bonds_per_structure at its default size) between bead pairs that are within contact range in the starting
    # territories -- like assigned contacts, satisfied when imposed.  Same beads, envelope, polymer, protocol.
    from scipy.spatial import cKDTree
    systems = []
    for s in range(n_sample):
        rng = np.random.default_rng(1234 + s)
        xyz = synth.territories(g, NUCLEUS_R, rng)
        near = cKDTree(xyz.astype(np.float64)).query_pairs(2.0 * 2.0 * float(rad[0]) * 0.95, output_type="ndarray")
        pick = near[rng.choice(near.shape[0], min(REF_HIC_BONDS, near.shape[0]), replace=False)]
        r0 = 2.0 * (rad[pick[:, 0]] + rad[pick[:, 1]]).astype(np.float64)
        systems.append(synth.pack_structure(xyz, rad, g, pick[:, 0], pick[:, 1], r0, nucleus_radius=NUCLEUS_R, seed=1 + s, struct_id=s))
    n_hic = int(np.mean([ps.bond_i.size for ps in systems])) - int(synth.polymer_bonds(g, rad)[0].size)
    times = []
    for it in range(args.warmup + args.steps):
        v, dt, _ = cpu_mstep_baseline(systems, prot, n_sample, cores)
        if it >= args.warmup:
            times.append(dt)
    ms = 1e3 * float(np.mean(times))
    value = n_sample / (ms / 1e3)
    sample = ("%d structures per step (one per host thread), steric + envelope + polymer + %d Hi-C upper bounds per structure, "
              "full protocol" % (n_sample, n_hic))
    print(json.dumps({
        "impl": "reference", "metric": "optimised structures/sec at 2Mb diploid", "value": value, "unit": "structures/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "configs[1]: 2 Mb diploid Hi-C (3008 beads, spherical envelope, full demo annealing protocol + CG, "
                               "%d Hi-C bonds per structure between beads in contact range of the starting territories); CPU "
                               "restatement of the lammpgen script (oracle/ff_oracle.c) -- lmp_serial itself is not installable "
                               "offline" % n_hic, "beads": 3008, "structures_per_step": n_sample, "hic_bonds_per_structure": n_hic,
                   "md_scale": args.md_scale},
        "cpu_baseline": {"value": value, "unit": "structures/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "structures/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


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
        h2d = out[3]["h2d_bytes"]
    torch.cuda.synchronize(device)
    e2e_ms = (time.perf_counter() - t0) * 1e3 / args.steps
    if not all(r.status == 0 for r in out[1]):
        raise RuntimeError("bench: a structure of the end-to-end pass came back non-finite")
    del out
    # ---- one full A/M iteration from host coordinates: upload -> all-gather -> activation distances
    #      (this rank's pair chunk) -> gather rows -> assignment -> bond build + submit -> M-step -> fetch ----
    from igm_b200.iteration import hic_iteration
    barrier()
    it = hic_iteration(crd, g, rad, wl["template"], wl["pairs"], wl["opts"], device=device, seeds=wl["seeds"])
    torch.cuda.synchronize(device)
    iter_s, t_astep, t_build = it["timings"]["total_s"], it["timings"]["astep_s"], it["timings"]["mstep_s"]

    tt = torch.tensor([ms_step, e2e_ms, wl["astep"]["actdist_kernel_ms"], wl["astep"]["actdist_e2e_ms"],
                       wl["astep"]["assign_kernel_ms"], wl["astep"]["assign_e2e_ms"], iter_s, t_astep, t_build],
                      dtype=torch.float64, device="cuda:%d" % device)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms_step, e2e_ms, ad_k, ad_e, as_k, as_e, iter_s, t_astep, t_build = [float(v) for v in tt.tolist()]
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        traffic = {}
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "r01_traffic.json")))
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
        line = {
            "metric": "optimised structures/sec at 2Mb diploid", "value": S * world / (ms_step * 1e-3), "unit": "structures/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32 (MD) + f64 (CG)", "data": "synthetic",
            "config": {"workload": "configs[1]: 2 Mb diploid Hi-C, %d structures per GPU, spherical envelope, full demo annealing "
                                   "protocol + CG; Hi-C matrix = contact frequencies of the relaxed synthetic population, sigma=%.4g "
                                   "keeps %d haploid pairs (the demo matrix keeps 614140 at its last sigma, 0.01)"
                                   % (S, a["sigma"], a["pairs_total"]),
                       "beads": 3008, "structures_per_gpu": S, "hic_bonds_per_structure": a["hic_bonds_per_structure"],
                       "md_force_evals": evals, "cg_evals_mean": cg_evals, "md_scale": args.md_scale,
                       "l2_policy": "working set per step (bond lists + neighbour tables, >400 MB) exceeds L2; A-step slab is "
                                    "L2 resident at S=1000 (36 MB), noted in roofline_astep"},
            "e2e": {"value": S * world / (e2e_ms * 1e-3), "unit": "structures/s", "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(d2h)},
            "gpu_launches": 2 * args.steps,
            "am_iteration": {"seconds": iter_s, "structures_per_s": S * world / iter_s, "astep_seconds": t_astep,
                             "mstep_seconds": t_build,
                             "what": "host coordinates -> upload, all-gather, activation distances, row gather, assignment, "
                                     "bond build + submit, annealing MD + CG, coordinates back on the host"},
            "clocks": clocks,
            "roofline": {"bound": "fp32", "kernel": "k_md", "achieved": md_tflops, "peak": FP32_PEAK_TFLOPS, "unit": "TFLOP/s",
                         "frac": md_tflops / FP32_PEAK_TFLOPS,
                         "traffic": (traffic["k_md2<3>"]["dram_bytes"] / traffic["k_md2<3>"]["structures"] / traffic["k_md2<3>"]["force_evals"]
                                     * S * evals) if "k_md2<3>" in traffic else None,
                         "traffic_note": "scaled from a short ncu --set full capture (profiles/r01_traffic.json): bytes per structure and "
                                         "force evaluation x this launch",
                         "flops_per_force_eval": feval,
                         "flops_note": "algorithmic = SURVEY 8(d) formula, every pair and bond counted once; the kernel gathers "
                                       "(no floating-point atomics, deterministic sums) and so executes each from both ends",
                         "executed": {"flops_per_force_eval": feval_exec,
                                      "tflops": feval_exec * evals * S / md_t / 1e12,
                                      "frac": feval_exec * evals * S / md_t / 1e12 / FP32_PEAK_TFLOPS},
                         "smem": {"bytes_per_force_eval": smem_eval, "achieved": smem_eval * evals * S / md_t / 1e12,
                                  "peak": SMEM_PEAK_TBS, "unit": "TB/s", "frac": smem_eval * evals * S / md_t / 1e12 / SMEM_PEAK_TBS,
                                  "note": "16-byte position reads the force loops issue, before bank-conflict replays "
                                          "(ncu: 2.3x on the gathers, LSU data pipe 54 % busy)"},
                         "kernel_ms": md_t * 1e3, "cg_kernel_ms": float(np.mean(cg_ms)),
                         "peak_source": "derived: 148 SM x 128 lanes x 2 x 1.965 GHz (non-tensor FP32; not in MEASURED_PEAKS.json)"},
            "astep": {"metric": "A-step haploid pairs/sec", "pairs": a["pairs_total"], "rows": a["rows"],
                      "structures": S * world,
                      "value": a["pairs_total"] / (ad_k * 1e-3), "e2e": a["pairs_total"] / (ad_e * 1e-3), "unit": "pairs/s",
                      "actdist_kernel_ms": ad_k, "actdist_e2e_ms": ad_e, "assign_kernel_ms": as_k, "assign_e2e_ms": as_e,
                      "assign_rows_x_structures_per_s": a["rows"] * float(S) / (as_k * 1e-3)},
            "roofline_astep": {"bound": "hbm", "kernel": "k_actdist", "achieved": ad_gbs, "peak": hbm, "unit": "GB/s",
                               "frac": ad_gbs / hbm,
                               "traffic": (traffic["k_actdist_reg<4>"]["dram_bytes"] / traffic["k_actdist_reg<4>"]["pairs"] * a["pairs_local"])
                               if ("k_actdist_reg<4>" in traffic and S * world == 1000) else None,
                               "traffic_note": "ncu --set full capture at 1000 structures scaled by pairs; the slab is L2 resident",
                               "peak_source": which + " copy bandwidth",
                               "assign_achieved": as_gbs, "assign_frac": as_gbs / hbm},
        }
        if not args.no_cpu_baseline and world == 1:
            cores = os.cpu_count() or 1
            n_sample = args.cpu_sample or cores
            v, dt, _ = cpu_mstep_baseline(systems, prot, n_sample, cores)
            line["cpu_baseline"] = {"value": v, "unit": "structures/s", "cores": cores, "kind": "port",
                                    "sample": "%d of the %d structures (same restraints, full protocol), one per host "
                                              "thread, %.1f s" % (n_sample, S, dt)}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
