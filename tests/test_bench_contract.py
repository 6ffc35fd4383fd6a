"""The bench line's contract (one JSON line; keys the driver and the judge read), checked on the committed round-2 lines and,
live, on the reference arm (which needs no GPU)."""
import glob
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE = ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
        "dtype", "data", "config")


def _check_line(d, ours=True):
    for k in BASE:
        assert k in d, k
    assert d["dtype"] == "f64" and d["data"] == "synthetic" and isinstance(d["config"].get("workload"), str)
    assert d["vs_baseline"] is None                       # BASELINE.md holds no published number for this metric
    if ours:
        r = d["roofline"]
        for k in ("bound", "achieved", "peak", "unit", "frac", "traffic"):
            assert k in r, k
        assert r["bound"] == "hbm" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
        assert d["gpu_launches"] > 0
        assert "clocks" in d


def test_committed_round2_lines_keep_the_contract():
    d = json.load(open(os.path.join(ROOT, "profiles", "r02_bench_final.json")))
    _check_line(d)
    assert d["metric"] == "simple_cell_updates_per_s" and d["unit"] == "GLUPS" and d["n_gpus"] == 1
    assert d["roofline"]["traffic"] is not None and "k_tdma_flow" in d["roofline"]["kernel"]
    for k in ("value", "unit", "cores", "kind", "sample"):
        assert k in d["cpu_baseline"], k
    for k in ("value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"):
        assert k in d["e2e"], k
    assert d["e2e"]["h2d_bytes_per_step"] > 4.0e8        # u, v, p of 4098^2 doubles really cross PCIe every step
    assert d["clocks"]["samples"] > 10 and not set(d["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    n = 0
    for f in glob.glob(os.path.join(ROOT, "profiles", "r02_scaling_final", "*.json")):
        d = json.load(open(f))
        _check_line(d)
        if d["metric"] == "simple_cell_updates_per_s" and d["n_gpus"] > 1:
            pc = d["parity_check"]
            assert pc["ok"] and pc["bitwise"] and pc["world"] == d["n_gpus"], f
            n += 1
    assert n >= 8


def test_reference_arm_prints_one_json_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "40", "--warmup", "1"],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-500:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    _check_line(d, ours=False)
    assert d["impl"] == "reference" and d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] == 1
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["value"] == d["value"]
    assert d["config"]["grid_timed"] == "1024x1024"       # 41 passes would not fit a few minutes at 4096^2: bounded sample
