"""bench.py's reference arm (the CPU side the driver runs as `--impl reference`) on a tiny workload: one JSON line on
stdout with the contract's keys, the honest-threading fields, and the keyed extras (Kernel SHAP port, configs 1 and 5)."""

import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(extra, env=None):
    cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "2", "--warmup", "1",
           "--cpu-rows", "40000", "--kshap-cpu-rows", "8", "--serving-calls", "5", *extra]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env={**os.environ, **(env or {})}, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, r.stdout[:500]              # stdout carries the one JSON line and nothing else
    return json.loads(lines[0])


def test_reference_arm_line():
    d = _run([])
    assert d["impl"] == "reference" and d["metric"] == "scored_rows_per_sec" and d["unit"] == "rows/s"
    assert d["higher_is_better"] is True and d["value"] > 0 and d["steps"] == 2 and d["vs_baseline"] is None
    assert d["config"]["workload"].startswith("config2") and d["config"]["features"] == 30
    cb = d["cpu_baseline"]
    assert cb["value"] == d["value"] and cb["cores"] >= 1 and cb["kind"] in ("reference", "port") and cb["sample"]
    assert set(cb["variants"]) == {"library_default", "all_cores"} and cb["variant"] in cb["variants"]
    assert cb["variants"]["all_cores"]["blas_threads"] == 1 and cb["variants"]["all_cores"]["python_threads"] == cb["cores"]
    assert d["e2e"] == {"value": d["value"], "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    ks = d["kernel_shap"]
    assert ks["unit"] == "explained rows/s" and ks["per_row_sets"]["value"] > 0 and ks["shared_set"]["value"] > 0
    assert d["config1"]["value"] > 0 and len(d["config1"]["top3"]) == 3
    assert {"batch_1", "batch_64", "batch_4096"} <= set(d["config5"])
    assert d["config5"]["batch_1"]["predict_bracket"]["p50_ms"] > 0


def test_non_zero_ranks_of_the_reference_arm_stay_silent():
    cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "1"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=120, cwd=ROOT,
                       env={**os.environ, "RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert r.returncode == 0 and r.stdout.strip() == ""
