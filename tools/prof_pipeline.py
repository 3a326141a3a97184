#!/usr/bin/env python
"""Wall time of kernel.run_population on the bench workload for several chunk sizes, and of
iteration.hic_iteration (diagnostics; not a bench value).  usage: prof_pipeline.py [--md-scale F] chunk ..."""
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench                                                     # noqa: E402


def main():
    argv = sys.argv[1:]
    md_scale = 1.0
    if argv and argv[0] == "--md-scale":
        md_scale = float(argv[1]); argv = argv[2:]
    chunks = [int(a) for a in argv] or [128]
    sys.argv = [sys.argv[0], "--md-scale", str(md_scale)]
    args = bench.parse_args()
    import torch
    from igm_b200.kernel import run_population
    from igm_b200.iteration import hic_iteration
    torch.cuda.set_device(0)
    wl = bench.build_workload(args, 0, 1, 0, lambda m: print("[wl] " + m, file=sys.stderr, flush=True))
    g, crd = wl["genome"], wl["crd"]
    row, col, dst, r0 = wl["rows"]
    for c in chunks:
        t0 = time.perf_counter()
        print("run_population starts at %.1f ms (steady clock)" % (time.monotonic() * 1e3), file=sys.stderr, flush=True)
        out = run_population(wl["template"], crd, len(g), row, col, r0, 1.0, wl["bits"], wl["prot"], seeds=wl["seeds"], chunk=c)
        dt = time.perf_counter() - t0
        print("run_population chunk=%d: %.3f s, sum of chunk md %.1f ms, cg %.1f ms" % (c, dt, out[3]["md_kernel_ms"], out[3]["cg_kernel_ms"]), flush=True)
    for _ in range(2):
        it = hic_iteration(crd, g, wl["radii"], wl["template"], wl["pairs"], wl["opts"], seeds=wl["seeds"])
        print("hic_iteration:", {k: round(v, 3) for k, v in it["timings"].items()}, flush=True)


if __name__ == "__main__":
    main()
