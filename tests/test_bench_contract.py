"""bench.py's JSON contract, checked where no GPU is needed: the reference arm (`--impl reference`) runs the reference's own
CPU implementation of the path — oracle/_ref (compiled from the reference's sources) for the configs it can take, the OpenMP
restatement for the rest — and prints ONE line with the same metric / unit / config as the GPU arm."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype", "data",
             "config", "cpu_baseline", "e2e", "impl"}


def run_ref(*extra):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", *extra], capture_output=True, text=True, cwd=ROOT,
                         timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, "exactly one JSON line on stdout"
    return json.loads(lines[0])


def check_common(d, steps, warmup):
    assert BASE_KEYS <= set(d), BASE_KEYS - set(d)
    assert d["impl"] == "reference" and d["metric"] == "particle_rk2_updates_per_s" and d["unit"] == "particle-updates/s"
    assert d["higher_is_better"] is True and d["vs_baseline"] is None and d["data"] == "synthetic" and d["n_gpus"] == 1
    assert d["steps"] == steps and d["warmup"] == warmup and d["value"] > 0 and d["ms_per_step"] > 0
    assert "workload" in d["config"] and "model" not in d["config"]
    cb, e2e = d["cpu_baseline"], d["e2e"]
    assert cb["value"] == d["value"] and cb["unit"] == d["unit"] and cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["sample"]
    assert e2e == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


@pytest.mark.skipif(not __import__("oracle").have_ref(), reason="compiled reference not built here")
def test_reference_arm_runs_the_compiled_reference_on_config_1():
    d = run_ref("--config", "c1", "--steps", "2", "--warmup", "1")
    check_common(d, 2, 1)
    assert d["cpu_baseline"]["kind"] == "reference" and d["cpu_baseline"]["cores"] == 1      # the reference has no threading
    assert "8x25x8" in d["config"]["workload"] and "10000 particles" in d["config"]["workload"]
    assert "FIT in L2" in d["config"]["l2"]                                                     # a parity-size case, says so


def test_reference_arm_falls_back_to_the_port_beyond_the_reference_ceiling():
    # 128x256x128 cannot run in the reference (SparseLU assembly is O(n^2)); bounded sample: few particles, 1 step
    d = run_ref("--config", "c3", "--particles", "50000", "--steps", "1", "--warmup", "0")
    check_common(d, 1, 0)
    assert d["cpu_baseline"]["kind"] == "port" and "128x256x128" in d["config"]["workload"]
    assert d["mac_cell_steps_per_s"] > 0
