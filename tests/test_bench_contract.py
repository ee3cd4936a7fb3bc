"""bench.py's JSON contract, checked on the arm that runs without a GPU (--impl reference)."""
import json
import subprocess
import sys

from conftest import ROOT, have_reference_binaries


def test_reference_arm_prints_one_json_line_with_the_contract_keys():
    pr = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--workload", "c1", "--steps", "20",
                         "--warmup", "3"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert pr.returncode == 0, pr.stderr[-2000:]
    lines = [l for l in pr.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "impl", "cpu_baseline", "e2e"):
        assert k in d, k
    assert d["impl"] == "reference" and d["unit"] == "GLUP/s" and d["higher_is_better"] is True
    assert d["vs_baseline"] is None and d["data"] == "synthetic" and "workload" in d["config"]
    assert d["value"] > 0 and d["ms_per_step"] > 0
    cb = d["cpu_baseline"]
    assert cb["value"] == d["value"] and cb["cores"] >= 1 and cb["sample"]
    assert cb["kind"] == ("reference" if have_reference_binaries() else "port")
    assert d["e2e"] == {"value": d["value"], "unit": "GLUP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_other_ranks_of_the_reference_arm_exit_quietly():
    import os
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    pr = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "5"],
                        capture_output=True, text=True, timeout=120, cwd=ROOT, env=env)
    assert pr.returncode == 0 and pr.stdout.strip() == ""


def test_bench_inidat_window_is_the_oracles_inidat(oracle):
    # bench.py checks windows of the timed field against the oracle started from inidat restricted to the window
    import importlib.util
    import numpy as np
    spec = importlib.util.spec_from_file_location("bench_mod", ROOT / "bench.py")
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    full = oracle.inidat(700, 900)  # wraps int32 beyond 432
    w = bench.inidat_window(700, 900, 430, 470, 500, 640)
    assert w.dtype == np.float32 and np.array_equal(w.view(np.uint32), full[430:470, 500:640].view(np.uint32))
    w64 = bench.inidat_window(700, 900, 430, 470, 500, 640, np.float64)
    assert np.array_equal(w64, oracle.inidat_f64(700, 900)[430:470, 500:640])
    assert bench.reference_grid(bench.workload("strong", 1))[:2] == (16384, 16384)
    assert bench.reference_grid(bench.workload("c2", 1))[2] is None
