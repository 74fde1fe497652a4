"""bench.py contract, CPU side: the reference arm prints ONE JSON line with the keys the driver reads; the product arm
refuses to run without a GPU (no CPU fallback).  The GPU side of the contract is exercised by tools/gpu_round.sh."""
import json
import os
import subprocess
import sys

import pytest
import torch as th

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(*args, env=None):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True,
                          cwd=ROOT, timeout=600, env=dict(os.environ, **(env or {})))


def test_reference_arm_line():
    p = run("--impl", "reference", "--steps", "1", "--warmup", "1")
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, lines
    line = json.loads(lines[0])
    assert line["impl"] == "reference" and line["higher_is_better"] is True and line["n_gpus"] == 1
    assert line["unit"] == "Gcandidates/s" and line["value"] > 0 and line["vs_baseline"] is None
    assert line["config"]["name"] == "c2" and "workload" in line["config"]
    cb = line["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == line["value"] and cb["sample"]
    assert line["e2e"] == {"value": line["value"], "unit": line["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_exit_quietly():
    p = run("--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "1",
            env=dict(RANK="1", LOCAL_RANK="1", WORLD_SIZE="2"))
    assert p.returncode == 0 and p.stdout.strip() == ""


@pytest.mark.skipif(th.cuda.is_available(), reason="needs a box without a GPU")
def test_product_arm_needs_a_gpu():
    p = run("--steps", "1", "--warmup", "1")
    assert p.returncode != 0 and "no CUDA device" in (p.stderr + p.stdout)
