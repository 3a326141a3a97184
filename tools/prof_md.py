#!/usr/bin/env python
"""M-step kernel timing on the bench fixture (tests/golden/bench_c2_structures.npz: M-step inputs of the first 32
structures of the default bench workload), replicated with different velocity seeds to fill the GPU.  Tuning aid:

    [IGMK_PROF=1] [IGMK_SKIN=..] [IGMK_SKIN_OUTER=..] python tools/prof_md.py [n_structures] [md_scale] [repeats]
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench                                              # noqa: E402
from igm_b200 import synth                                # noqa: E402
from igm_b200.kernel import MStepBatch                    # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 148
    md_scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
    reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    fx = bench.load_fixture()
    g = synth.genome_2mb()
    rad = synth.bead_radii(len(g), bench.NUCLEUS_R, 0.4)
    base = bench.fixture_systems(fx, g, rad, bench.N_FIXTURE)
    _, prot = bench.protocol(md_scale)
    evals = bench.md_force_evals(prot)
    with MStepBatch(0) as b:
        for s in range(n):
            ps = base[s % len(base)]
            ps.c.seed = 7 + s
            b.submit(ps)
        t0 = time.perf_counter()
        b.upload()
        t_up = time.perf_counter() - t0
        out = []
        for _ in range(reps):
            b.reset_coordinates()
            b.run(prot)
            out.append(b.last_timings()[:2])
        e = np.array([b.fetch(s)[1].final_energy for s in range(n)])
        rb = np.array([b.fetch(s)[1].n_rebuilds for s in range(n)])
    md = min(o[0] for o in out)
    waves = max(1.0, n / 148.0)
    print("structures %d, force evals %d: md %.1f ms (%.2f us per force evaluation and wave), cg %.1f ms, upload %.2f s; "
          "median final energy %.4g, finite %d/%d, rebuilds %.0f"
          % (n, evals, md, md * 1e3 / evals / waves, out[-1][1], t_up, np.median(e), int(np.isfinite(e).sum()), n, rb.mean()))


if __name__ == "__main__":
    main()
