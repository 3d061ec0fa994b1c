"""bench.py as the round's driver launches it, as far as that goes without a GPU: the reference arm (the reference's own
place.cpp on the host cores, oracle/_ref) prints the contract's line, under a multi-rank launch only rank 0 does, and
the product arm refuses to run without a CUDA device (there is no CPU path to fall back to)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BENCH = os.path.join(ROOT, "bench.py")


def _run(argv, **env):
    return subprocess.run([sys.executable, BENCH] + argv, capture_output=True, text=True, cwd=ROOT,
                          env=dict(os.environ, EPK_BENCH_REF_SECONDS="1", **env), timeout=600)


def test_reference_arm_prints_the_contract_line(built):
    run = _run(["--impl", "reference", "--gpus", "1", "--steps", "1", "--warmup", "1"])
    assert run.returncode == 0, run.stderr[-2000:]
    lines = [l for l in run.stdout.split("\n") if l.startswith("{")]
    assert len(lines) == 1
    line = json.loads(lines[0])
    assert line["impl"] == "reference" and line["metric"] == "placed queries/sec" and line["unit"] == "reads/s"
    assert line["higher_is_better"] is True and line["n_gpus"] == 1 and line["steps"] == 1 and line["warmup"] == 1
    assert line["value"] > 0 and line["ms_per_step"] > 0 and line["vs_baseline"] is None
    assert line["config"]["workload"].startswith("C2")
    base = line["cpu_baseline"]
    assert base["kind"] in ("reference", "port") and base["cores"] >= 1 and base["value"] == line["value"] and "reads" in base["sample"]
    assert line["e2e"] == {"value": line["value"], "unit": "reads/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_runs_on_rank_zero_only():
    run = _run(["--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"], RANK="1", LOCAL_RANK="1", WORLD_SIZE="2")
    assert run.returncode == 0 and run.stdout.strip() == "", run.stdout + run.stderr[-2000:]


def test_product_arm_needs_a_cuda_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    run = _run(["--steps", "1", "--warmup", "0"])
    assert run.returncode != 0 and "no CUDA device" in (run.stderr + run.stdout)
