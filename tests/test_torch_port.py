"""The torch-CPU port used as the timed CPU baseline must compute what the
reference computes: pinned against the same golden vectors (CPU only)."""
import numpy as np
import pytest
import torch

from oracle import torch_port as TP
from oracle import vae_numpy as O
from tests.golden_cases import PARAM_SEED, batch_seed, check_fp, load


@pytest.mark.parametrize("tag", ["step_b128", "step_b48"])
def test_port_matches_reference_golden(tag):
    g = load(tag)
    B, n = int(g["B"]), int(g["n_steps"])
    torch.set_num_threads(4)
    model = TP.VAEPort()
    TP.load_segments(model, O.init_params(PARAM_SEED))
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    for t in range(n):
        x, eps = O.make_batch(batch_seed(t), B)
        loss = TP.train_step(model, opt, torch.from_numpy(x).view(B, 1, 28, 28), torch.from_numpy(eps))
        assert abs(loss - g["losses"][t]) <= 2e-6 * abs(g["losses"][t])
    seg = TP.to_segments(model)
    for k in O.SEGMENTS:
        check_fp(g, "paramN", k, seg[k], 1e-5, tag)


def test_port_timed_steps_counts():
    model = TP.VAEPort()
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    batches = [torch.rand(8, 1, 28, 28) for _ in range(2)]
    n, dt = TP.timed_steps(model, opt, batches, warmup=1, steps=5, seconds=None)
    assert n == 5 and dt > 0
    n, dt = TP.timed_steps(model, opt, batches, warmup=0, steps=None, seconds=0.05)
    assert n >= 1 and dt >= 0.05
