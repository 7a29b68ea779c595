"""bench.py's contract on a machine without a GPU: the reference arm runs (it is the CPU oracle) and prints ONE JSON line
with the driver's keys; the product arm refuses to run without a CUDA device instead of falling back to anything."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*flags):
    env = dict(os.environ)
    env.pop("RANK", None)
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *flags], capture_output=True, text=True,
                          timeout=600, cwd=ROOT, env=env)


def test_reference_arm_prints_one_json_line_with_the_contract_keys():
    p = _run("--impl", "reference", "--steps", "1", "--warmup", "0", "--cpu-sample", "4096", "--no-extra")
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    rec = json.loads(lines[0])
    assert rec["impl"] == "reference" and rec["metric"] == "logdensity_samples_per_s" and rec["unit"] == "samples/s"
    assert rec["higher_is_better"] is True and rec["value"] > 0 and rec["steps"] == 1 and rec["gpu_launches"] == 0
    cb = rec["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == rec["value"] and "sample" in cb
    assert rec["e2e"] == {"value": rec["value"], "unit": rec["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert rec["config"]["workload"].startswith("C2") and rec["config"]["sample_per_step"] == 4096


def test_reference_arm_under_torchrun_only_rank_zero_works():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                       capture_output=True, text=True, timeout=120, cwd=ROOT, env=env)
    assert p.returncode == 0 and p.stdout.strip() == ""


def test_product_arm_refuses_to_run_without_a_cuda_device():
    import torch
    if torch.cuda.is_available():
        import pytest
        pytest.skip("a GPU is present")
    p = _run("--steps", "1", "--warmup", "0")
    assert p.returncode != 0 and "no CPU fallback" in (p.stderr + p.stdout)
