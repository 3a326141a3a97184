#!/usr/bin/env python
"""Timing of the large-structure M-step path at BASELINE config 4 size (hg38 male at 200 kb, 29 838 beads):
   python tools/prof_c4.py [n_struct] [md_scale]
territories -> relaxation pass -> the same structures again from the relaxed coordinates with ~20 k
sequence-local upper bounds each (stand-in for assigned Hi-C restraints)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from igm_b200 import synth, _abi                 # noqa: E402
from igm_b200.kernel import MStepBatch           # noqa: E402

S = int(sys.argv[1]) if len(sys.argv) > 1 else 4
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 0.05
g = synth.genome_200kb()
R = 5500.0
rad = synth.bead_radii(len(g), R, 0.4)
opts = synth.demo_protocol_options()
cap = opts["custom_annealing_protocol"]
cap["mdsteps"] = [max(1, int(v * scale)) for v in cap["mdsteps"]]
cap["relax"]["mdsteps"] = max(1, int(cap["relax"]["mdsteps"] * scale))
prot = _abi.make_protocol(opts)
evals = sum(prot.mdsteps[k] + 1 + prot.relax_mdsteps + 1 for k in range(prot.n_stage))
start = [synth.territories(g, R, np.random.default_rng(50 + s)) for s in range(S)]
with MStepBatch(0) as b:
    for s in range(S):
        b.submit(synth.pack_structure(start[s], rad, g, nucleus_radius=R, seed=1 + s, struct_id=s))
    b.run(prot)
    relaxed = [b.fetch(s)[0][:len(g)] for s in range(S)]
    print("relaxation pass: %.0f ms MD (%d force evaluations) + %.0f ms CG for %d structures" % (b.last_timings()[0], evals, b.last_timings()[1], S))
rng = np.random.default_rng(3)
with MStepBatch(0) as b:
    for s in range(S):
        hi = rng.integers(0, len(g), 20000)
        hj = np.minimum(hi + rng.integers(2, 40, 20000), len(g) - 1)
        b.submit(synth.pack_structure(relaxed[s], rad, g, hi, hj, np.full(20000, 4.0 * float(rad[0])), nucleus_radius=R, seed=9 + s, struct_id=s))
    b.run(prot)
    md_ms, cg_ms, _ = b.last_timings()
    res = b.fetch(0)[1]
    print("restrained run: %.0f ms MD = %.0f us per force evaluation and structure-CTA (%d rebuilds), CG %.0f ms (%d evaluations); "
          "full protocol (47 008 evaluations) ~ %.0f s per structure" % (md_ms, 1e3 * md_ms / evals, res.n_rebuilds, cg_ms, res.cg_evals, md_ms / evals * 47008 / 1e3))
