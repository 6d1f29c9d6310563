"""CPU: bench.py's reference arm prints the contract's JSON line on a test-scale workload; the product arm refuses to run
without a CUDA device (no CPU fallback); workload definitions are consistent."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "t1", "--steps", "2", "--warmup", "1"],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["metric"] == "Mrays/s" and line["unit"] == "Mrays/s" and line["higher_is_better"] is True
    assert line["value"] > 0 and line["steps"] == 2 and line["warmup"] == 1 and line["gpu_launches"] == 0
    assert line["cpu_baseline"]["kind"] in ("reference", "port") and line["cpu_baseline"]["cores"] >= 1
    assert line["cpu_baseline"]["value"] == line["value"] == line["e2e"]["value"]
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
    assert "workload" in line["config"] and "model" not in line["config"]


def test_reference_arm_nonzero_ranks_exit_quietly():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "t1", "--gpus", "2", "--steps", "1"],
                         capture_output=True, text=True, timeout=300, cwd=ROOT, env=env)
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_product_arm_needs_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--workload", "t1", "--steps", "1", "--warmup", "0"],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode != 0 and "no CPU fallback" in (out.stderr + out.stdout)


def test_workloads_and_counts():
    from blok_b200 import workloads
    for name in ("c2", "c3", "c4", "t1"):
        w = workloads.WORKLOADS[name]
        assert w["kind"] in ("primary", "primary_shadow", "path") and w["width"] % 8 == 0 and w["height"] % 8 == 0
    assert workloads.rays_per_step("c3") == 3840 * 2160
    c0, c1 = workloads.camera_for("c3", 0), workloads.camera_for("c3", 1)
    assert c1.position[0] == c0.position[0] + 16.0 and c1.position[1:] == c0.position[1:]
    with open(os.path.join(ROOT, "blok_b200", "workload_counts.json")) as f:
        cc = json.load(f)
    assert cc["c3"]["V"] > 1 and cc["c3"]["A"] > 0.1
