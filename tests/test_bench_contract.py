"""bench.py's JSON contract, checked on the committed round-1 lines (profiles/) and on the
argument parser; no GPU needed. The lines themselves come from runs on a B200 (see profiles/)."""

import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE_KEYS = ["metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better",
             "scaling", "vs_baseline", "dtype", "data", "config", "roofline", "cpu_baseline", "e2e",
             "clocks", "gpu_launches"]


def _load(name):
    path = os.path.join(ROOT, "profiles", name)
    if not os.path.exists(path):
        pytest.skip(f"{name} not recorded")
    return json.load(open(path))


def test_own_arm_line():
    d = _load("r01_bench_n1_final.json")
    for k in BASE_KEYS:
        assert k in d, k
    assert d["n_gpus"] == 1 and d["higher_is_better"] is True and d["unit"] == "tokens/s"
    assert d["vs_baseline"] is None  # BASELINE.md has no published number for this exact metric
    assert "workload" in d["config"] and "model" not in d["config"]
    r = d["roofline"]
    for k in ("bound", "achieved", "peak", "unit", "frac", "traffic"):
        assert k in r, k
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    c = d["cpu_baseline"]
    for k in ("value", "unit", "cores", "kind", "sample"):
        assert k in c, k
    assert c["kind"] in ("port", "reference")
    e = d["e2e"]
    for k in ("value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"):
        assert k in e, k
    assert e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0 and e["value"] < d["value"]
    assert d["gpu_launches"] > 0
    assert {"sm_mhz", "sm_max_mhz", "reasons"} <= set(d["clocks"])
    assert not ({"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"} & set(d["clocks"]["reasons"]))
    # the >= 70 % of the nominal 8 TB/s HBM roofline bar of north_star: 6.14 M tok/s at V = 151,936
    assert d["value"] >= 6.14e6


def test_reference_arm_line():
    d = _load("r01_bench_reference_arm.json")
    assert d["impl"] == "reference" and d["unit"] == "tokens/s"
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["cpu_baseline"]["value"] == d["value"] and d["cpu_baseline"]["cores"] >= 1


def test_scaling_lines_are_whole_job_aggregates():
    one = _load("r01_bench_n1_final.json")["value"]
    for n in (2, 4, 8):
        d = _load(f"r01_bench_n{n}.json")
        assert d["n_gpus"] == n and d["scaling"] == "weak"
        assert 0.85 * n * one < d["value"] < 1.1 * n * one  # total over all ranks, not per GPU


def test_bench_cli_parses_without_a_gpu():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--help"], capture_output=True, text=True)
    assert r.returncode == 0
    for flag in ("--gpus", "--steps", "--warmup", "--impl"):
        assert flag in r.stdout
