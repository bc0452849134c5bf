"""bench.py's reference arm (CPU cores, runs anywhere) prints the contract's JSON line -- also when it is started the
way the driver starts it for N > 1 (python -m torch.distributed.run, whose TORCHELASTIC_* variables must not leak into
the CPU workers) -- and the own arm refuses to run without a GPU instead of falling back to anything."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(args, timeout=400, launcher=None):
    cmd = [sys.executable] + (launcher or []) + [os.path.join(ROOT, "bench.py")] + args
    return subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=timeout)


def check_line(r, n):
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip().startswith("{")]
    assert len(lines) == 1, r.stdout
    j = json.loads(lines[0])
    assert j["impl"] == "reference" and j["metric"] == "VAE-HPO samples/sec (whole box)" and j["unit"] == "samples/s"
    assert j["higher_is_better"] is True and j["n_gpus"] == n and j["value"] > 0 and j["vs_baseline"] is None
    assert j["dtype"] == "f32" and j["data"] == "synthetic"
    assert set(j["config"]) == {"workload", "layout", "batch_per_gpu", "hidden"} and j["config"]["layout"] == f"{n}x1"
    cb = j["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["value"] == j["value"] and cb["sample"]
    assert j["e2e"] == {"value": j["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    return j


def test_reference_arm_prints_one_contract_line():
    j = check_line(run(["--impl", "reference", "--steps", "2", "--warmup", "1"]), 1)
    # in the build container the reference itself is what runs (oracle/_ref or /root/reference)
    if os.path.exists("/root/reference/vae-hpo.py") or os.path.exists(os.path.join(ROOT, "oracle", "_ref", "vae-hpo.py")):
        assert j["cpu_baseline"]["kind"] == "reference"


def test_reference_arm_under_torchrun_with_two_ranks():
    """The driver's N = 2 launch.  Round 1's arm hung here for 828 s: torchrun exports TORCHELASTIC_USE_AGENT_STORE=True,
    which made every gloo worker a client of a TCPStore nobody served."""
    launcher = ["-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                "--master-port", "29791"]
    check_line(run(["--impl", "reference", "--gpus", "2", "--steps", "3", "--warmup", "1"], timeout=240, launcher=launcher), 2)


def test_both_arms_emit_the_same_config_dict():
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    b = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(b)
    src = open(os.path.join(ROOT, "bench.py")).read()
    # every "config" the file emits (both arms, and the watchdog's partial line) comes from config_of(): nothing arm-specific
    assert src.count('"config": config_of(') >= 2 and src.count('"config":') == src.count('"config": config_of(')
    assert b.config_of(8, 128, 400)["layout"] == "8x1"


def test_own_arm_needs_a_gpu_and_says_so():
    import torch
    if torch.cuda.is_available():
        return          # on a GPU box the own arm is exercised by the driver and by profiles/
    r = run(["--steps", "2", "--warmup", "1"], timeout=200)
    assert r.returncode != 0
    assert "CUDA" in (r.stderr + r.stdout) or "GPU" in (r.stderr + r.stdout)
