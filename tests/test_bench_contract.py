"""bench.py's CPU-runnable parts: the reference arm prints one JSON line with the contract's keys, and the
product arm refuses to run without a GPU (no CPU fallback)."""
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_json_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "1"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "TFLOP/s" and d["higher_is_better"] is True
    assert d["config"]["workload"] == "case4_gspmd_ff" and d["config"]["d_ff"] == 16384
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert d["e2e"]["value"] == d["value"] > 0
    assert d["vs_baseline"] is None
    # the config object is built by the same function in both arms: the driver's same-config comparison holds
    sys.path.insert(0, ROOT)
    import bench
    assert d["config"] == bench.workload_config("case4_gspmd_ff", bench.WORKLOADS["case4_gspmd_ff"], "gelu_tanh", 1)


def test_reference_arm_other_ranks_exit_quietly():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
                          "--steps", "1", "--warmup", "1"], capture_output=True, text=True, timeout=120, cwd=ROOT, env=env)
    assert out.returncode == 0 and out.stdout.strip() == ""


@pytest.mark.skipif(torch.cuda.is_available(), reason="only meaningful on a GPU-less box")
def test_product_arm_needs_a_gpu():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1", "--warmup", "3"],
                         capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert out.returncode != 0
    assert "no CPU fallback" in (out.stderr + out.stdout)


def test_bench_parity_bound_is_the_tests_bound():
    """bench.py restates the tolerance of tests/_tol.py (so that it needs nothing from tests/ at run time): same number."""
    sys.path.insert(0, ROOT)
    import bench
    from tests import _tol
    g = torch.Generator().manual_seed(0)
    want = torch.randn(64, 128, generator=g) * 0.05
    got = want + torch.randn(64, 128, generator=g) * 2e-4
    need, ulps, max_abs = bench._scaled_need(got, want)
    assert abs(need - _tol.scaled_error(got, want, 2e-2)) < 1e-9
    assert max_abs == float((got.double() - want.double()).abs().max()) and ulps > 0
