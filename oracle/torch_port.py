"""torch-CPU port of the reference's trial step (TEST / BASELINE INFRASTRUCTURE ONLY).

The reference is Python and cannot travel to the GPU box, so the timed CPU
baseline (`bench.py --impl reference` and the `cpu_baseline` leg) runs this
restatement: the same torch calls the reference makes, on the host cores, through
the same `utils`-style group bootstrap and real `DistributedDataParallel` over gloo.

* model            ref:vae-hpo.py:19-45   (nn.Linear x5, relu, exp, randn_like, sigmoid)
* loss             ref:vae-hpo.py:49-58   (F.binary_cross_entropy(sum) + KLD)
* step             ref:vae-hpo.py:69-74   (zero_grad, forward, loss, backward, loss.item(), step)
* trial set-up     ref:vae-hpo.py:129-131 (DDP(model, process_group=group), Adam(lr=1e-3))

Parameters live under the reference's names (fc1 ... fc4), so state dicts are
interchangeable with the reference's VAE; tests/test_torch_port.py pins this port
against the same golden vectors as the numpy oracle.  Nothing under
multidisttorch_b200/ imports this file.
"""
from __future__ import annotations

import time

import torch
import torch.nn.functional as F
from torch import nn


class VAEPort(nn.Module):
    def __init__(self, in_dim=784, hidden=400, latent=20):
        super().__init__()
        self.in_dim = in_dim
        self.fc1 = nn.Linear(in_dim, hidden)
        self.fc21 = nn.Linear(hidden, latent)
        self.fc22 = nn.Linear(hidden, latent)
        self.fc3 = nn.Linear(latent, hidden)
        self.fc4 = nn.Linear(hidden, in_dim)

    def decode(self, z):
        return torch.sigmoid(self.fc4(F.relu(self.fc3(z))))

    def forward(self, x, eps=None):
        h1 = F.relu(self.fc1(x.view(-1, self.in_dim)))
        mu, logvar = self.fc21(h1), self.fc22(h1)
        std = torch.exp(0.5 * logvar)
        if eps is None:
            eps = torch.randn_like(std)
        return self.decode(mu + eps * std), mu, logvar


def loss_port(recon, x, mu, logvar, in_dim=784):
    bce = F.binary_cross_entropy(recon, x.view(-1, in_dim), reduction="sum")
    return bce - 0.5 * torch.sum(1 + logvar - mu.pow(2) - logvar.exp())


def load_segments(model: VAEPort, seg: dict):
    """oracle.vae_numpy segment dict -> model parameters."""
    Z = seg["W2x"].shape[0] // 2
    t = lambda a: torch.as_tensor(a).clone()
    sd = {"fc1.weight": t(seg["W1"]), "fc1.bias": t(seg["b1"]),
          "fc21.weight": t(seg["W2x"][:Z]), "fc21.bias": t(seg["b2x"][:Z]),
          "fc22.weight": t(seg["W2x"][Z:]), "fc22.bias": t(seg["b2x"][Z:]),
          "fc3.weight": t(seg["W3"]), "fc3.bias": t(seg["b3"]),
          "fc4.weight": t(seg["W4"]), "fc4.bias": t(seg["b4"])}
    model.load_state_dict(sd)


def to_segments(model: VAEPort) -> dict:
    sd = {k: v.detach().cpu().numpy() for k, v in model.state_dict().items()}
    import numpy as np
    return {"W1": sd["fc1.weight"], "b1": sd["fc1.bias"],
            "W2x": np.concatenate([sd["fc21.weight"], sd["fc22.weight"]], 0),
            "b2x": np.concatenate([sd["fc21.bias"], sd["fc22.bias"]], 0),
            "W3": sd["fc3.weight"], "b3": sd["fc3.bias"], "W4": sd["fc4.weight"], "b4": sd["fc4.bias"]}


def train_step(model, opt, x, eps=None, in_dim=784):
    """ref:vae-hpo.py:69-74, including the per-step loss.item() host read."""
    opt.zero_grad()
    recon, mu, logvar = model(x, eps) if eps is not None else model(x)
    loss = loss_port(recon, x, mu, logvar, in_dim)
    loss.backward()
    val = loss.item()
    opt.step()
    return val


def timed_steps(model, opt, batches, warmup, steps=None, seconds=None, barrier=None):
    """Run `warmup` untimed steps then either exactly `steps` or as many as fit in
    `seconds`; returns (n_steps, elapsed_seconds)."""
    n = len(batches)
    for i in range(warmup):
        train_step(model, opt, batches[i % n])
    if barrier:
        barrier()
    t0 = time.perf_counter()
    done = 0
    while True:
        train_step(model, opt, batches[done % n])
        done += 1
        if steps is not None and done >= steps:
            break
        if steps is None and time.perf_counter() - t0 >= seconds:
            break
    if barrier:
        barrier()
    return done, time.perf_counter() - t0
