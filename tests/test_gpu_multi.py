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
    logs = [p.communicate(timeout=600)[0] for p in procs]
    assert [p.returncode for p in procs] == [0] * world, "\n".join(l[-1500:] for l in logs)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("mode", ["fp32", "tf32x3"])
def test_one_group_of_two_gpus_matches_reference_ddp(mode):
    run_world(2, "ddp_1x2_b32", mode)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_one_group_of_two_gpus_nccl_path_matches_reference_ddp():
    run_world(2, "ddp_1x2_b32", "tf32x3", {"MDT_P2P_ALLREDUCE": "0"})


@pytest.mark.skipif(torch.cuda.device_count() < 4, reason="needs 4 GPUs")
@pytest.mark.parametrize("mode", ["fp32", "tf32x3"])
def test_two_groups_of_two_gpus_match_reference_ddp(mode):
    run_world(4, "ddp_2x2_b16", mode)
