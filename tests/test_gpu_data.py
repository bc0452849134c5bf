"""Device-resident data path against the reference's own host pipeline:
datasets.MNIST + ToTensor + DistributedSampler(rank=group_id, num_replicas=ngroups) + DataLoader
(ref:vae-hpo.py:135-150) on synthetic idx files."""
import struct

import numpy as np
import pytest
import torch

from oracle import vae_numpy as O

pytestmark = pytest.mark.gpu


def _write_idx(path, arr):
    with open(path, "wb") as f:
        if arr.ndim == 3:
            f.write(struct.pack(">IIII", 0x00000803, *arr.shape))
        else:
            f.write(struct.pack(">II", 0x00000801, arr.shape[0]))
        f.write(arr.tobytes())


def test_device_batches_equal_dataloader_batches_and_training_matches(tmp_path):
    from torchvision import datasets, transforms
    from multidisttorch_b200 import dropin
    from multidisttorch_b200.data import DeviceDataset
    from multidisttorch_b200.engine import VAEEngine
    dropin.uninstall()          # this test compares against torch's OWN DataLoader (an earlier `import utils` may have
                                # put the device-resident loader in its place)

    raw = tmp_path / "MNIST" / "raw"
    raw.mkdir(parents=True)
    rs = np.random.RandomState(3)
    n = 1000                                            # 1000 / 2 groups = 500 per group -> 3 full batches + one of 116
    images = rs.randint(0, 256, (n, 28, 28), dtype=np.uint8)
    for split in ("train", "t10k"):
        _write_idx(raw / f"{split}-images-idx3-ubyte", images)
        _write_idx(raw / f"{split}-labels-idx1-ubyte", rs.randint(0, 10, (n,), dtype=np.uint8))
    trainset = datasets.MNIST(str(tmp_path), train=True, download=True, transform=transforms.ToTensor())
    group_id, ngroups, B = 1, 2, 128
    sampler = torch.utils.data.distributed.DistributedSampler(trainset, rank=group_id, num_replicas=ngroups)
    loader = torch.utils.data.DataLoader(trainset, batch_size=B, shuffle=False, sampler=sampler)
    host_batches = [data for data, _ in loader]
    assert [b.shape[0] for b in host_batches] == [128, 128, 128, 116]

    ds = DeviceDataset(images)
    idx = ds.sampler_indices(rank=group_id, num_replicas=ngroups)
    assert idx.cpu().tolist() == list(iter(sampler))
    for i, hb in enumerate(host_batches):
        out = torch.empty(hb.shape[0], 784, device="cuda")
        ds.gather(idx[i * B:(i + 1) * B], out)
        assert torch.equal(out.cpu(), hb.view(-1, 784))          # bit-exact: uint8 -> float32 / 255

    # the same epoch, host-fed vs device-resident, ends with identical losses and parameters
    p0 = O.init_params(9)
    a, b = VAEEngine(max_batch=B), VAEEngine(max_batch=B)
    for e in (a, b):
        e.load_segments(p0)
        e.set_rng(77, 0)
    la = [float(a.train_step(hb.view(-1, 784).cuda(), None).item()) for hb in host_batches]
    lb = [float(l.item()) for l in ds.epoch(b, idx, B)]
    assert la == lb
    assert torch.equal(a.params, b.params)
    a.close(); b.close()
