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


def test_round2_own_arm_line():
    d = _load("r02_bench_n1_final.json")
    for k in BASE_KEYS + ["roofline_gae", "strong_scaling", "parity"]:
        assert k in d, k
    assert d["n_gpus"] == 1 and d["parity"] is None  # the cross-GPU checks need N >= 2
    assert d["value"] >= 6.14e6  # north_star: >= 70 % of the nominal 8 TB/s roofline at 1 GPU
    r = d["roofline"]
    assert r["bound"] == "hbm" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9 and r["frac"] > 0.9
    assert r["traffic"] and 0.9 < r["traffic"] / r["algorithmic_bytes_per_launch"] < 1.05  # no wasted re-reads
    # the CPU leg ran the reference's own functions (baseline/_ref travelled to the box)
    c = d["cpu_baseline"]
    assert c["kind"] == "reference" and c["cores"] >= 1 and "4096-token" in c["sample"]
    # GAE has its own roofline entries, with the kernel named and the CPU baseline beside cfg1
    g = d["roofline_gae"]
    assert len(g) == 4 and all(x["bound"] == "hbm" and x["algorithmic_bytes_per_token"] == 16 for x in g)
    assert any(x["cpu_baseline_tokens_per_s"] for x in g)
    # forward-only and fp32-logits call sites are in the line
    v = d["extra"]["h1_call_site_variants"]
    assert {x["logits"] for x in v} == {"bfloat16", "float32"} and all(x["frac_of_measured_peak"] > 0.85 for x in v)
    # three launches per micro-batch on the drop-in path
    assert d["gpu_launches"] == d["steps"] * (3 + 3 * d["config"]["micro_batches"])


def test_round2_reference_arm_line():
    d = _load("r02_bench_reference_arm.json")
    own = _load("r02_bench_n1_final.json")
    assert d["impl"] == "reference" and d["cpu_baseline"]["kind"] == "reference"
    assert d["metric"] == own["metric"] and d["unit"] == own["unit"] and d["config"]["workload"] == own["config"]["workload"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    # both legs time the same sample size, so the two CPU numbers agree to a few per cent
    assert abs(d["value"] / own["cpu_baseline"]["value"] - 1) < 0.15


@pytest.mark.parametrize("n", [2, 4, 8])
def test_round2_multi_gpu_lines_carry_parity(n):
    d = _load(f"r02_bench_n{n}.json")
    assert d["n_gpus"] == n and d["scaling"] == "weak"
    p = d["parity"]
    assert p and all(p[k] is True for k in ("sharded_step", "group_norm", "tp", "peer_move"))
    det = p["details"]
    assert det["tp"]["tp"] == n and det["tp"]["max_abs_err"]["grad_ulp_viol"] == 0
    assert det["peer_move"]["consumers"] == n - 1
    e = det["sharded_step"]["err"]
    assert e["adv_max_abs"] <= 1e-5 and e["loss_rel"] <= 1e-3 and e["g_logp_rel"] <= 1e-3 and e["total_loss_weight"] == 0.0
    s = d["strong_scaling"]
    assert s["scaling"] == "strong" and s["imbalance_max_over_mean_tokens"] <= 1.01


def test_round2_summary_matches_the_committed_lines():
    """profiles/r02_SUMMARY.md is generated from the committed bench lines; it must not drift."""
    import importlib.util

    spec = importlib.util.spec_from_file_location("summarize_bench", os.path.join(ROOT, "scripts", "summarize_bench.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    with open(os.path.join(ROOT, "profiles", "r02_SUMMARY.md")) as f:
        assert f.read() == mod.build()
