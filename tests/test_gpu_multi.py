"""Multi-GPU parity (needs >= 2 GPUs; skipped on a 1-GPU box): per-group gradient exchange -- over NVLink peer memory
inside the Adam kernel (default on the tensor-core path) and with ncclAllReduce on the subgroup's communicator --
checked against the golden vectors of the reference's real DDP runs."""
import os
import subprocess
import sys
import tempfile

import pytest
import torch

from tests.dist_launch import next_port

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def run_world(world, tag, mode, extra_env=None):
    out = tempfile.mkdtemp(prefix="mdt_gpu_multi_")
    port = next_port()
    procs = []
    for r in range(world):
        env = dict(os.environ)
        for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "DDP_BACKEND"):
            env.pop(k, None)
        env.update(OMPI_COMM_WORLD_SIZE=str(world), OMPI_COMM_WORLD_RANK=str(r), OMPI_COMM_WORLD_LOCAL_RANK=str(r),
                   MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), MDT_DROPIN="0",
                   PYTHONPATH=os.path.dirname(HERE) + os.pathsep + env.get("PYTHONPATH", ""))
        env.update(extra_env or {})
        procs.append(subprocess.Popen([sys.executable, os.path.join(HERE, "_gpu_ddp_worker.py"), tag, mode, out], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    # fail fast: the moment one rank exits non-zero its peers (blocked in a collective or a barrier) are killed, and the
    # whole world gets 240 s -- a hang must not hold N GPUs for long
    import time
    t0 = time.time()
    while time.time() - t0 < 240:
        codes = [p.poll() for p in procs]
        if all(c is not None for c in codes) or any(c not in (None, 0) for c in codes):
            break
        time.sleep(0.5)
    time.sleep(1.0)
    for p in procs:
        if p.poll() is None:
            p.kill()
    logs = [p.communicate()[0] for p in procs]
    assert [p.returncode for p in procs] == [0] * world, "\n".join(f"--- rank {r} (rc {p.returncode})\n{l[-1500:]}" for r, (p, l) in enumerate(zip(procs, logs)))


def need(n):
    return pytest.mark.skipif(torch.cuda.device_count() < n, reason=f"needs {n} GPUs")


@need(2)
@pytest.mark.parametrize("mode", ["fp32", "tf32x3", "bf16"])
def test_one_group_of_two_gpus_matches_reference_ddp(mode):
    run_world(2, "ddp_1x2_b32", mode)


@need(2)
@pytest.mark.parametrize("xch", ["nccl", "p2p", "p2p_two_shot"])
def test_one_group_of_two_gpus_every_exchange_matches_reference_ddp(xch):
    run_world(2, "ddp_1x2_b32", "tf32x3", {"MDT_EXCHANGE": xch})


@need(4)
@pytest.mark.parametrize("mode", ["fp32", "tf32x3"])
def test_two_groups_of_two_gpus_match_reference_ddp(mode):
    run_world(4, "ddp_2x2_b16", mode)


@need(4)
@pytest.mark.parametrize("xch", ["nccl", "p2p", "p2p_two_shot"])
def test_one_group_of_four_gpus_matches_reference_ddp(xch):
    run_world(4, "ddp_1x4_b32", "tf32x3", {"MDT_EXCHANGE": xch})


@need(8)
@pytest.mark.parametrize("tag", ["ddp_1x8_b16", "ddp_2x4_b128"])
@pytest.mark.parametrize("xch", ["nccl", "p2p", "p2p_two_shot"])
def test_eight_gpu_layouts_match_reference_ddp(tag, xch):
    """1x8 (BASELINE config 5's layout) and 2x4 at 128 / rank (config 4's layout and batch)."""
    run_world(8, tag, "tf32x3", {"MDT_EXCHANGE": xch})
