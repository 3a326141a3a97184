#!/usr/bin/env python
"""Does the demo comparison (tests/test_gpu_demo.py) pin the bond energy prefactor?  Runs the demo schedule down to
sigma 0.02 with E_bond = k dr^2 (prefactor 1, the LAMMPS `harmonic` convention assumed throughout) and with
E_bond = k/2 dr^2 (prefactor 0.5, the IGM docstrings' reading) and prints, for each, the statistics that test asserts
next to the reference's recorded population and the test's tolerances.

    python tools/demo_prefactor.py [workdir]      (inputs: oracle/_ref/demo/, staged by oracle/build_ref.sh)"""
import json
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import popstats                                                     # noqa: E402
from igm_b200 import files, synth                                   # noqa: E402
from igm_b200.iteration import hic_iteration_files, relax_init      # noqa: E402

DEMO = os.path.join(ROOT, "oracle", "_ref", "demo")


def run(pref, work):
    start, hcs = os.path.join(DEMO, "igm-model.hss.randomInit"), os.path.join(DEMO, "WTC11_HiC_2Mb.hcs")
    opts = synth.demo_protocol_options()
    pop = files.read_hss(start)
    idx, rad, S = pop["index"], pop["radii"], pop["nstruct"]
    R = float(np.asarray(pop["tree"]["envelope"]["params"]).ravel()[0])
    template = synth.pack_structure(np.zeros((idx.n_bead, 3), dtype=np.float32), rad, idx, nucleus_radius=R, seed=0, struct_id=0)
    crd, _ = relax_init(pop["coordinates"], template, opts, seeds=np.arange(S, dtype=np.int64) + 1, bond_prefactor=pref)
    cur = os.path.join(work, "relaxed_%g.hss" % pref)
    files.write_hss(cur, pop["tree"], coordinates=crd)
    last, out = None, None
    for step_no, sigma in enumerate([1.0, 0.2, 0.1, 0.05, 0.02], start=1):
        nxt, ad = os.path.join(work, "s%d_%g.hss" % (step_no, pref)), os.path.join(work, "a%d_%g.hdf5" % (step_no, pref))
        out = hic_iteration_files(cur, hcs, nxt, ad, opts, sigma, sigma, last_actdist=last, bond_prefactor=pref,
                                  seeds=np.arange(S, dtype=np.int64) + 1 + 1000 * step_no, plast_override=1.0 - 1e-7)
        cur, last = nxt, ad
    return files.read_hss(cur), out["summary"]


def main():
    work = sys.argv[1] if len(sys.argv) > 1 else tempfile.mkdtemp()
    os.makedirs(work, exist_ok=True)
    ref = files.read_hss(os.path.join(DEMO, "igm-model.hss.T"))
    rs = json.loads(ref["tree"]["summary"])
    rb = {k.split("[")[0]: v["n_imposed"] for k, v in rs["byrestraint"].items()}
    b = dict(popstats.structure_stats(ref))
    keys = (["P(contact), separation %d" % s for s in (1, 2, 3, 5, 10, 20)] +
            ["mean distance / r, separation %d" % s for s in (1, 2, 3, 5, 10, 20)] +
            ["mean radial position / R", "radius of gyration of chr1 copy 0 / r"])
    tol = {k: 0.02 for k in keys if k.startswith("P(")}
    tol.update({k: 0.05 * b[k] for k in keys if k.startswith("mean distance")})
    tol["mean radial position / R"] = 0.005
    tol["radius of gyration of chr1 copy 0 / r"] = 0.4
    res = {}
    for pref in (1.0, 0.5):
        mine, ms = run(pref, work)
        a = dict(popstats.structure_stats(mine))
        res[pref] = (a, ms)
    print("%-46s %10s %12s %12s %10s" % ("statistic", "reference", "prefactor 1", "prefactor 0.5", "tolerance"))
    hic_ref = rb["interHiC"] + rb["intraHiC"]
    print("%-46s %10d %12d %12d %10s" % ("Hi-C restraints imposed", hic_ref, res[1.0][1]["byrestraint"]["Polymer+HiC"]["n_imposed_hic"],
                                       res[0.5][1]["byrestraint"]["Polymer+HiC"]["n_imposed_hic"], "0.1 %"))
    print("%-46s %10d %12d %12d %10s" % ("violations (> 5 %)", rs["n_violations"], res[1.0][1]["n_violations"], res[0.5][1]["n_violations"], "<= 10"))
    outside = []
    for k in keys:
        d1, d5 = res[1.0][0][k] - b[k], res[0.5][0][k] - b[k]
        flag = "  <-- outside" if abs(d5) >= tol[k] else ""
        if flag:
            outside.append(k)
        print("%-46s %10.4f %+12.4f %+12.4f %10.4f%s" % (k, b[k], d1, d5, tol[k], flag))
    hic5 = res[0.5][1]["byrestraint"]["Polymer+HiC"]["n_imposed_hic"]
    if abs(hic5 / hic_ref - 1.0) >= 1e-3:
        outside.append("Hi-C restraints imposed")
    if res[0.5][1]["n_violations"] > 10:
        outside.append("violations")
    print("statistics of the prefactor-0.5 run outside the tolerances of tests/test_gpu_demo.py: %s" % (outside or "none"))


if __name__ == "__main__":
    main()
