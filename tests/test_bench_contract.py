"""The measurement contract of bench.py: the reference arm runs here on the host cores and prints one JSON line with
the contract's keys; the committed final line of the GPU arm (profiles/) carries the same keys plus roofline,
cpu_baseline, e2e with copies, clocks and the launch count."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE_KEYS = ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
             "vs_baseline", "dtype", "data", "config")


def test_reference_arm_prints_the_contract_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "small",
                          "--steps", "1", "--warmup", "0"], capture_output=True, text=True, cwd=ROOT, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    d = json.loads(out.stdout.strip().splitlines()[-1])
    for k in BASE_KEYS:
        assert k in d, k
    assert d["impl"] == "reference" and d["higher_is_better"] is True and d["value"] > 0
    assert d["metric"] == "cell-layer updates/s" and "workload" in d["config"]
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["sample"] and cb["value"] == d["value"]
    e = d["e2e"]
    assert e["value"] == d["value"] and e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0


def test_committed_gpu_line_carries_every_contract_key():
    path = os.path.join(ROOT, "profiles", "r02u_bench_config3_default_run.json")
    d = json.loads(open(path).read().strip().splitlines()[-1])
    for k in BASE_KEYS:
        assert k in d, k
    base = json.load(open(os.path.join(ROOT, "BASELINE.json")))
    assert d["metric"] == base.get("metric", d["metric"]) or d["metric"] in str(base)
    assert d["dtype"] == "f64" and d["data"] == "synthetic" and d["n_gpus"] == 1 and d["warmup"] >= 3
    assert "655362" in d["config"]["workload"].replace(" ", "") or "655 362" in d["config"]["workload"]
    r = d["roofline"]
    for k in ("bound", "achieved", "peak", "unit", "frac", "traffic"):
        assert k in r, k
    assert r["bound"] == "hbm" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    e = d["e2e"]
    assert e["h2d_bytes_per_step"] > 4e9 and e["d2h_bytes_per_step"] > 4e9 and e["value"] < d["value"]
    cb = d["cpu_baseline"]
    for k in ("value", "unit", "cores", "kind", "sample"):
        assert k in cb, k
    c = d["clocks"]
    assert c["sm_mhz"] > 0.9 * c["sm_max_mhz"]
    assert not set(c["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    assert d["gpu_launches"] == 16 * d["steps"]
