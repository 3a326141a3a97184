"""bench.py's reference arm runs on the CPU: check the JSON contract of that line here (the GPU arm
prints the same keys plus roofline / astep objects and is exercised on the GPU box)."""
import json
import os
import subprocess
import sys

from conftest import ROOT


def test_reference_arm_json_line():
    out = subprocess.check_output([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                                   "--warmup", "0", "--md-scale", "0.002", "--cpu-sample", "2"], cwd=ROOT, timeout=300)
    lines = [l for l in out.decode().splitlines() if l.strip()]
    assert len(lines) == 1
    j = json.loads(lines[0])
    if not os.path.isfile(os.path.join(ROOT, "tests", "golden", "bench_c2_structures.npz")):
        assert j["impl"] == "reference" and "unavailable" in j        # the fixture is made on a B200 (bench.py --write-fixture)
        return
    for key in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert key in j, key
    assert j["impl"] == "reference" and j["unit"] == "structures/s" and j["higher_is_better"] is True
    assert j["cpu_baseline"]["kind"] == "port" and j["cpu_baseline"]["cores"] >= 1 and j["cpu_baseline"]["value"] == j["value"]
    assert j["e2e"] == {"value": j["value"], "unit": j["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert j["value"] > 0 and "workload" in j["config"]
    # both arms print the same `config`: it is built from the arguments alone
    sys.path.insert(0, ROOT)
    import bench
    import argparse
    args = argparse.Namespace(structures=1000, pairs=614140, md_scale=0.002)
    assert j["config"] == bench.workload_config(args, j["config"]["md_force_evals"])
    assert j["workload_stats"]["structures_per_step"] == 2


def test_reference_arm_other_ranks_do_nothing():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    out = subprocess.check_output([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                                  cwd=ROOT, env=env, timeout=120)
    assert out.decode().strip() == ""
