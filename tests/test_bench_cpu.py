"""bench.py's reference arm (the one leg that runs without a GPU): one JSON line with the contract's keys for every
workload name, rank 0 only under a multi-rank launch."""
import json
import os
import subprocess
import sys

import pytest

from _oracle import Ref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = {"impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
        "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"}


def run(args, env=None):
    e = dict(os.environ); e.update(env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, capture_output=True, text=True, env=e, timeout=600)


@pytest.mark.skipif(not Ref.available(), reason="oracle/_ref/libcvref.so not built")
@pytest.mark.parametrize("workload,shape", [("1080p", "1920x1080"), ("vga", "640x480"), ("kitti", "1242x375")])
def test_reference_arm_line(workload, shape):
    p = run(["--impl", "reference", "--steps", "1", "--warmup", "0", "--workload", workload])
    assert p.returncode == 0, p.stderr
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert KEYS <= set(d) and d["impl"] == "reference" and d["unit"] == "frames/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["e2e"] == {"value": d["value"], "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["cpu_baseline"]["kind"] == "reference" and d["cpu_baseline"]["cores"] >= 1 and shape in d["cpu_baseline"]["sample"]
    assert "workload" in d["config"] and "model" not in d["config"]


def test_reference_arm_other_ranks_print_nothing():
    p = run(["--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"], env={"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert p.returncode == 0 and p.stdout.strip() == ""


def test_product_arm_refuses_to_run_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    p = run(["--steps", "1", "--warmup", "1"])
    assert p.returncode != 0 and "no CPU fallback" in (p.stderr + p.stdout)
