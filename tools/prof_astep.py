#!/usr/bin/env python
"""Small driver for profiling the A-step kernels alone (ncu target):
   python tools/prof_astep.py [n_struct] [n_pairs]
Population = synthetic territories (no relaxation), Hi-C matrix = its own contact frequencies."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from igm_b200 import synth                     # noqa: E402
from igm_b200.astep import Population          # noqa: E402

S = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
NP = int(sys.argv[2]) if len(sys.argv) > 2 else 614140
g = synth.genome_2mb()
rad = synth.bead_radii(len(g))
crd = synth.population_territories(g, S, seed=99)
with Population(S, rad, g.chrom, g.copy_off, g.copy_beads) as pop:
    pop.upload(crd)
    iu, ju = np.triu_indices(g.n_hap, 1)
    freq = pop.contact_freq(iu.astype(np.int32), ju.astype(np.int32), 2.0)
    pw = np.minimum(np.float32(1), np.float32(1.1) * freq)
    thr = np.partition(pw, pw.size - NP)[pw.size - NP]
    keep = pw >= thr
    pi, pj = iu[keep].astype(np.int32), ju[keep].astype(np.int32)
    pl = np.round(freq[keep].astype(np.float64), 4)
    for rep in range(3):
        t = time.perf_counter()
        rows = pop.actdist(pi, pj, pw[keep].astype(np.float64), pl, 2.0)
        e2e = time.perf_counter() - t
        k_ad = pop.last_kernel_ms()
    bits = pop.assign(rows[0], rows[1], rows[2])
    k_as = pop.last_kernel_ms()
    nc = np.diff(g.copy_off)
    nbytes = 12.0 * S * (nc[pi] + nc[pj]).sum() + 32.0 * pi.size
    print("S=%d pairs=%d rows=%d actdist kernel %.3f ms (%.1f Mpairs/s, %.0f GB/s algorithmic) e2e %.1f ms; assign %.3f ms"
          % (S, pi.size, rows[0].size, k_ad, pi.size / k_ad / 1e3, nbytes / k_ad / 1e6, e2e * 1e3, k_as))
