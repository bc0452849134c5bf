"""bench.py's reference arm (CPU cores, runs anywhere) prints the contract's JSON line; the own arm refuses to run
without a GPU instead of falling back to anything."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(args, timeout=400):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, cwd=ROOT, capture_output=True,
                          text=True, timeout=timeout)


def test_reference_arm_prints_one_contract_line():
    r = run(["--impl", "reference", "--steps", "2", "--warmup", "1", "--cpu-seconds", "2"])
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, r.stdout
    j = json.loads(lines[0])
    assert j["impl"] == "reference" and j["metric"] == "VAE-HPO samples/sec (whole box)" and j["unit"] == "samples/s"
    assert j["higher_is_better"] is True and j["n_gpus"] == 1 and j["value"] > 0 and j["vs_baseline"] is None
    assert j["dtype"] == "f32" and j["data"] == "synthetic" and "workload" in j["config"]
    cb = j["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == j["value"] and cb["sample"]
    assert j["e2e"] == {"value": j["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_own_arm_needs_a_gpu_and_says_so():
    import torch
    if torch.cuda.is_available():
        return          # on a GPU box the own arm is exercised by the driver and by profiles/
    r = run(["--steps", "2", "--warmup", "1"], timeout=200)
    assert r.returncode != 0
    assert "CUDA" in (r.stderr + r.stdout) or "GPU" in (r.stderr + r.stdout)
