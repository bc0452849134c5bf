"""The reference's own scripts, UNCHANGED, on top of this repo's `utils` drop-in
(CPU/gloo: BASELINE config 1 "plumbing, no GPU").  Needs /root/reference, which
exists only in the build container -- skipped elsewhere."""
import os
import struct
import sys

import numpy as np
import pytest

from tests import dist_launch

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.exists(os.path.join(REF, "vae-hpo.py")), reason="reference not mounted")


def test_example_subgroup_runs_unchanged():
    out, codes, logs = dist_launch.launch(8, [], script=os.path.join(REF, "example-subgroup.py"))
    assert codes == [0] * 8, logs
    for r, log in enumerate(logs):
        exp = "[tensor([0]), tensor([1]), tensor([2]), tensor([3])]" if r < 4 else "[tensor([4]), tensor([5]), tensor([6]), tensor([7])]"
        assert f"{r} gather_list: {exp}" in log, log
        assert f"DDP setup: 8 {r}" in log and "Done." in log


def _write_idx(path, arr):
    with open(path, "wb") as f:
        if arr.ndim == 3:
            f.write(struct.pack(">IIII", 0x00000803, *arr.shape))
        else:
            f.write(struct.pack(">II", 0x00000801, arr.shape[0]))
        f.write(arr.tobytes())


def test_vae_hpo_runs_unchanged_two_groups(tmp_path):
    raw = tmp_path / "data" / "MNIST" / "raw"
    raw.mkdir(parents=True)
    rs = np.random.RandomState(0)
    for split, n in (("train", 1024), ("t10k", 256)):
        _write_idx(raw / f"{split}-images-idx3-ubyte", rs.randint(0, 256, (n, 28, 28), dtype=np.uint8))
        _write_idx(raw / f"{split}-labels-idx1-ubyte", rs.randint(0, 10, (n,), dtype=np.uint8))
    out, codes, logs = dist_launch.launch(2, ["--ngroups", "2", "--epochs", "1", "--batch-size", "128"],
                                          script=os.path.join(REF, "vae-hpo.py"), cwd=str(tmp_path), timeout=600)
    assert codes == [0, 0], logs
    assert "Rank 0 is in group 0" in logs[0] and "Rank 1 is in group 1" in logs[1]
    assert "[0:0] ====> Epoch: 1 Average loss:" in logs[0] and "[1:0] ====> Test set loss:" in logs[1]
    assert "[1:0] Train Epoch: 2" in logs[1] and "Train Epoch: 2" not in logs[0]     # epochs + group_id (ref:vae-hpo.py:202)
    assert "0 Done. time:" in logs[0] and "1 Done. time:" in logs[1]
