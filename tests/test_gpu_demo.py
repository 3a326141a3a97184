"""north_star criterion 3 against the reference's OWN recorded run: the demo (BASELINE configs[0]) from the reference's
RandomInit file through RelaxInit and the A/M schedule down to sigma 0.02, file to file on the GPU, must land on the
population lammpgen produced in the reference's demo output (igm-model.hss.T, the population after the sigma = 0.02
modeling step).  The recorded run imposed p = pwish restraints per pair (tests/test_files.py), so the
schedule here is fed the same (plast just below 1 makes t = 0 and p = pwish; kernels and arithmetic unchanged).
Inputs: oracle/_ref/demo/ (data files of the reference, staged by oracle/build_ref.sh; skipped when absent).
Tolerances are stated at the assertions; measured differences are in profiles/r01_demo_vs_reference.txt."""
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
DEMO = os.path.join(ROOT, "oracle", "_ref", "demo")
FILES = [os.path.join(DEMO, n) for n in ("igm-model.hss.randomInit", "WTC11_HiC_2Mb.hcs", "igm-model.hss.T")]

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not all(os.path.isfile(f) for f in FILES), reason="oracle/_ref/demo not staged")]


def test_demo_population_matches_the_reference_recorded_output(tmp_path):
    import popstats
    from igm_b200 import files, synth
    from igm_b200.iteration import hic_iteration_files, relax_init
    start, hcs, recorded = FILES
    opts = synth.demo_protocol_options()
    pop = files.read_hss(start)
    idx, rad, S = pop["index"], pop["radii"], pop["nstruct"]
    R = float(np.asarray(pop["tree"]["envelope"]["params"]).ravel()[0])
    template = synth.pack_structure(np.zeros((idx.n_bead, 3), dtype=np.float32), rad, idx, nucleus_radius=R, seed=0, struct_id=0)
    crd, _ = relax_init(pop["coordinates"], template, opts, seeds=np.arange(S, dtype=np.int64) + 1)
    cur = str(tmp_path / "relaxed.hss")
    files.write_hss(cur, pop["tree"], coordinates=crd)
    last, out = None, None
    for step_no, sigma in enumerate([1.0, 0.2, 0.1, 0.05, 0.02], start=1):          # demo/config_file.json sigma lists
        nxt, ad = str(tmp_path / ("step%d.hss" % step_no)), str(tmp_path / ("actdist%d.hdf5" % step_no))
        out = hic_iteration_files(cur, hcs, nxt, ad, opts, sigma, sigma, last_actdist=last,
                                  seeds=np.arange(S, dtype=np.int64) + 1 + 1000 * step_no, plast_override=1.0 - 1e-7)
        s = out["summary"]
        assert s["n_violations"] < 0.01 * s["n_imposed"]                             # max_violations: the schedule moves on
        cur, last = nxt, ad
    mine, ref = files.read_hss(cur), files.read_hss(recorded)
    rs = json.loads(ref["tree"]["summary"])
    rb = {k.split("[")[0]: v["n_imposed"] for k, v in rs["byrestraint"].items()}
    ms = out["summary"]
    # restraint counts: envelope and polymer identical, Hi-C and the total within 0.1 % (measured: +0.004 %)
    assert ms["byrestraint"]["Envelope"]["n_imposed"] == rb["Envelope"]
    assert ms["byrestraint"]["Polymer+HiC"]["n_imposed_polymer"] == rb["Polymer"]
    hic_ref = rb["interHiC"] + rb["intraHiC"]
    assert abs(ms["byrestraint"]["Polymer+HiC"]["n_imposed_hic"] / hic_ref - 1.0) < 1e-3
    assert abs(ms["n_imposed"] / rs["n_imposed"] - 1.0) < 1e-3
    assert ms["n_violations"] <= 10                                                  # recorded: 3 of 1.44 M
    # structure of the populations (tools/popstats.py)
    a, b = dict(popstats.structure_stats(mine)), dict(popstats.structure_stats(ref))
    for sep in (1, 2, 3, 5, 10, 20):
        assert abs(a["P(contact), separation %d" % sep] - b["P(contact), separation %d" % sep]) < 0.02      # measured <= 0.005
        assert abs(a["mean distance / r, separation %d" % sep] - b["mean distance / r, separation %d" % sep]) < 0.05 * b["mean distance / r, separation %d" % sep]
    assert abs(a["mean radial position / R"] - b["mean radial position / R"]) < 0.005                       # measured 0.0001
    assert abs(a["radius of gyration of chr1 copy 0 / r"] - b["radius of gyration of chr1 copy 0 / r"]) < 0.4   # measured 0.08 of 7.4
    assert a["random bead pairs: P(d < 2r)"] < 1e-3                                                          # no overlaps left
