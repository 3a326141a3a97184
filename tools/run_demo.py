#!/usr/bin/env python
"""The reference's demo (BASELINE configs[0]: demo/config_file.json on demo/WTC11_HiC_2Mb.hcs, 100 structures, 2 Mb)
through the B200 path, file to file: RelaxInit, then the A/M schedule of bin/igm-run:87-260 -- sigma list 1.0 .. 0.01,
min_iterations 5, max_iterations 10, max_violations 0.01, force_last_iteration -- with the iteration control
restated here (it is the reference's control plane, not part of the product).

    python tools/run_demo.py <igm-model.hss.randomInit> <WTC11_HiC_2Mb.hcs> <workdir> [--expect igm-model.hss.T] [--md-scale F]

Prints one line per A/M iteration and, with --expect, the restraint statistics of the final population next to the
ones the reference recorded in its own demo output.  The input files are the reference's (copy them under
oracle/_ref/demo/ to take them to a GPU box: git-ignored, travels with gpurun)."""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from igm_b200 import files, synth                                   # noqa: E402
from igm_b200.iteration import hic_iteration_files, relax_init      # noqa: E402

SIGMAS = [1.0, 0.2, 0.1, 0.05, 0.02, 0.01]          # demo/config_file.json: intra_sigma_list == inter_sigma_list
MIN_ITER, MAX_ITER, MAX_VIOLATIONS = 5, 10, 0.01     # optimization/min_iterations, max_iterations, max_violations


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("hss"); ap.add_argument("hcs"); ap.add_argument("workdir")
    ap.add_argument("--expect", default=None, help="the reference's own final population file, for the comparison")
    ap.add_argument("--md-scale", type=float, default=1.0)
    ap.add_argument("--no-correction", action="store_true",
                    help="activation distances from p = pwish (the line commented out at ActivationDistanceStep.py:437) instead of "
                         "the corrected probability: plast = 1 - 1e-7 for every pair makes t = 0 and p = pwish")
    ap.add_argument("--stop-sigma", type=float, default=0.0, help="stop after the first acceptable iteration at this sigma")
    args = ap.parse_args()
    os.makedirs(args.workdir, exist_ok=True)
    opts = synth.demo_protocol_options()
    if args.md_scale != 1.0:
        cap = opts["custom_annealing_protocol"]
        cap["mdsteps"] = [max(1, int(v * args.md_scale)) for v in cap["mdsteps"]]
        cap["relax"]["mdsteps"] = max(1, int(cap["relax"]["mdsteps"] * args.md_scale))
    t_all = time.perf_counter()
    pop = files.read_hss(args.hss)
    idx, rad, S = pop["index"], pop["radii"], pop["nstruct"]
    R = float(np.asarray(pop["tree"]["envelope"]["params"]).ravel()[0])
    template = synth.pack_structure(np.zeros((idx.n_bead, 3), dtype=np.float32), rad, idx, nucleus_radius=R, seed=0, struct_id=0)
    t0 = time.perf_counter()
    crd, _ = relax_init(pop["coordinates"], template, opts, seeds=np.arange(S, dtype=np.int64) + 1)
    cur = os.path.join(args.workdir, "relaxed.hss")
    files.write_hss(cur, pop["tree"], coordinates=crd)
    print("RelaxInit: %d structures x %d beads in %.2f s" % (S, idx.n_bead, time.perf_counter() - t0), flush=True)

    last_ad, step_no, n_iter, t_am = None, 0, 0, 0.0
    for sigma in [s for s in SIGMAS if s >= args.stop_sigma]:
        opt_iter = 0
        while True:
            step_no += 1
            nxt = os.path.join(args.workdir, "step%02d.hss" % step_no)
            ad = os.path.join(args.workdir, "actdist%02d.hdf5" % step_no)
            t0 = time.perf_counter()
            out = hic_iteration_files(cur, args.hcs, nxt, ad, opts, sigma, sigma, last_actdist=last_ad,
                                      seeds=np.arange(S, dtype=np.int64) + 1 + 1000 * step_no,
                                      plast_override=(1.0 - 1e-7) if args.no_correction else None)
            dt = time.perf_counter() - t0
            t_am += dt; n_iter += 1
            s = out["summary"]
            score = s["n_violations"] / max(s["n_imposed"], 1)
            print("sigma %.3g iter %d: %d rows, %d restraints, %d violated (score %.3g); %.2f s (A-step %.3f s, M-step %.2f s)"
                  % (sigma, opt_iter, out["rows"][0].size, s["n_imposed"], s["n_violations"], score, dt,
                     out["timings"]["astep_s"], out["timings"]["mstep_s"]), flush=True)
            cur, last_ad = nxt, ad
            # bin/igm-run:137-260
            acceptable = score < MAX_VIOLATIONS
            force_next = (opt_iter < MIN_ITER - 1) and opt_iter != 0
            if sigma == SIGMAS[-1] and opt_iter == 0 and not args.stop_sigma:            # force_last_iteration: everything included, one iteration run
                force_next = True
            if acceptable and not force_next:
                break
            opt_iter += 1
            if opt_iter >= MAX_ITER:
                print("maximum number of iterations reached at sigma %g" % sigma)
                break
    print("schedule done: %d A/M iterations in %.1f s (%.2f s each, %.1f structures/s), %.1f s with file I/O and RelaxInit"
          % (n_iter, t_am, t_am / n_iter, S * n_iter / t_am, time.perf_counter() - t_all))
    final = files.read_hss(cur)
    mine = json.loads(final["tree"]["summary"])
    print("final population: %s" % json.dumps(mine))
    if args.expect:
        ref = json.loads(files.read_hss(args.expect)["tree"]["summary"])
        br = {k.split("[")[0]: v for k, v in ref["byrestraint"].items()}
        print("reference's demo output: n_imposed %d, n_violations %d; Envelope %d, Polymer %d, interHiC %d, intraHiC %d"
              % (ref["n_imposed"], ref["n_violations"], br["Envelope"]["n_imposed"], br["Polymer"]["n_imposed"],
                 br["interHiC"]["n_imposed"], br["intraHiC"]["n_imposed"]))
        hic_ref = br["interHiC"]["n_imposed"] + br["intraHiC"]["n_imposed"]
        hic_mine = mine["byrestraint"]["Polymer+HiC"]["n_imposed_hic"]
        print("Hi-C restraints imposed: %d here, %d in the reference's run (%.2f %%); violated fraction %.2e here, %.2e there"
              % (hic_mine, hic_ref, 100.0 * hic_mine / hic_ref, mine["n_violations"] / mine["n_imposed"],
                 ref["n_violations"] / ref["n_imposed"]))


if __name__ == "__main__":
    main()
