"""The reference's own vae-hpo.py, UNCHANGED, on a GPU: once on plain torch-CUDA (+ NCCL DDP) and once with this
repo's drop-in active (`from utils import *` resolves to the repo's utils.py, which installs the engine behind
DistributedDataParallel / optim.Adam).  Both runs start from the same generator seed (set by the runner, outside
the script); the drop-in draws eps from the same Philox stream torch.randn_like would have used -- the printed
losses of the two runs must agree.

The script is read from oracle/_ref/ (staged by oracle/vendor_ref.py from /root/reference; that directory travels to
the GPU box) and executed with runpy, so nothing in it is edited.  MNIST cannot be downloaded here: synthetic idx
files of the same format stand in (torchvision checks existence only)."""
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

from tests.dist_launch import next_port
from tests.test_reference_scripts import _write_idx

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCRIPT = os.path.join(ROOT, "oracle", "_ref", "vae-hpo.py")
# The script seeds nothing and torch seeds its default generators from the OS, so two runs of it never agree; the
# runner seeds torch (CPU and CUDA generators) BEFORE handing control to the untouched script.
RUNNER = ("import runpy, sys, torch; torch.manual_seed(1234); sys.argv = sys.argv[1:]; "
          "runpy.run_path(sys.argv[0], run_name='__main__')")


def make_data(tmp_path):
    raw = tmp_path / "data" / "MNIST" / "raw"
    raw.mkdir(parents=True)
    rs = np.random.RandomState(0)
    for split, n in (("train", 1024), ("t10k", 256)):
        _write_idx(raw / f"{split}-images-idx3-ubyte", rs.randint(0, 256, (n, 28, 28), dtype=np.uint8))
        _write_idx(raw / f"{split}-labels-idx1-ubyte", rs.randint(0, 10, (n,), dtype=np.uint8))


def run_script(tmp_path, world, ngroups, dropin):
    port = next_port()
    procs = []
    for r in range(world):
        env = dict(os.environ)
        for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "DDP_BACKEND"):
            env.pop(k, None)
        env.update(OMPI_COMM_WORLD_SIZE=str(world), OMPI_COMM_WORLD_RANK=str(r), OMPI_COMM_WORLD_LOCAL_RANK=str(r),
                   MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), MDT_DROPIN="1" if dropin else "0",
                   NVIDIA_TF32_OVERRIDE="0", PYTHONPATH=ROOT + os.pathsep + env.get("PYTHONPATH", ""))
        if not dropin:
            env["CUDA_VISIBLE_DEVICES"] = str(r)       # the reference uses the bare "cuda" device (ref:vae-hpo.py:127)
        procs.append(subprocess.Popen([sys.executable, "-c", RUNNER, SCRIPT, "--ngroups", str(ngroups), "--epochs", "1",
                                       "--batch-size", "128"], env=env, cwd=str(tmp_path), stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT, text=True))
    import time
    t0 = time.time()
    while time.time() - t0 < 300:
        codes = [p.poll() for p in procs]
        if all(c is not None for c in codes) or any(c not in (None, 0) for c in codes):
            break
        time.sleep(0.5)
    time.sleep(1.0)
    for p in procs:
        if p.poll() is None:
            p.kill()
    logs = [p.communicate()[0] for p in procs]
    assert [p.returncode for p in procs] == [0] * world, "\n".join(l[-3000:] for l in logs)
    return logs


def losses_of(log):
    return [float(v) for v in re.findall(r"(?:Loss|Average loss|Test set loss): ([0-9.]+)", log)]


@pytest.mark.skipif(not os.path.exists(SCRIPT), reason="oracle/_ref not staged (python oracle/vendor_ref.py)")
@pytest.mark.parametrize("world", [1, 2])
def test_unchanged_vae_hpo_through_the_dropin_matches_plain_torch_cuda(tmp_path, world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    make_data(tmp_path)
    plain = run_script(tmp_path, world, 1, dropin=False)
    ours = run_script(tmp_path, world, 1, dropin=True)
    assert "-> B200 engine" in ours[0], ours[0][-2000:]
    assert "device-resident loader" in ours[0], ours[0][-2000:]      # the script's MNIST DataLoaders were taken over too
    if world > 1:
        assert "gradient exchange: peer memory" in ours[0] or "gradient exchange: NCCL" in ours[0]
    a, b = losses_of(plain[0]), losses_of(ours[0])
    assert len(a) == len(b) and len(a) >= 3, (plain[0][-1500:], ours[0][-1500:])
    for u, v in zip(a, b):
        assert abs(u - v) <= 2e-4 * abs(u), (a, b)
    assert "0 Done. time:" in ours[0]
