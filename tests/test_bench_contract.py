"""bench.py's reference arm runs on the host cores only, so its JSON contract is checked without a GPU: one line,
the keys the driver reads, `impl: reference`, the oracle named as what was timed, no transfers."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_contract_line():
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                        "--ref-size", "10"], capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, lines          # stdout carries the JSON line and nothing else
    d = json.loads(lines[0])
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
                "dtype", "data", "config", "impl", "cpu_baseline", "e2e"):
        assert key in d, key
    assert d["impl"] == "reference" and d["n_gpus"] == 1 and d["steps"] == 1 and d["higher_is_better"] is True
    assert d["vs_baseline"] is None and d["dtype"] == "f64" and d["data"] == "synthetic"
    assert d["value"] > 0 and d["ms_per_step"] > 0
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] == 1 and cb["value"] == d["value"] and "10^3" in cb["sample"]
    e = d["e2e"]
    assert e["value"] == d["value"] and e["unit"] == d["unit"] and e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0
    # the sample is named for what it is: not the 256^3 inputs of the other arm
    assert d["config"]["same_config"] is False and "256^3" in d["config"]["workload"]


def test_committed_profiles_are_what_bench_reads():
    """the roofline's `traffic` comes from a committed ncu capture: the file bench.py names exists and carries the key"""
    src = open(os.path.join(ROOT, "bench.py")).read()
    name = "r02_ncu_mg_level_kernels_full.json"
    assert name in src
    pj = json.load(open(os.path.join(ROOT, "profiles", name)))
    assert pj["dram_bytes_per_launch_avg"] > 5e8 and len(pj["launches"]) >= 8
