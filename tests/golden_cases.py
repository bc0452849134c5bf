"""Shared helpers for reading tests/golden/* (outputs of the real reference,
written by oracle/make_golden.py) and regenerating the matching inputs."""
import os

import numpy as np

from oracle import vae_numpy as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
PARAM_SEED = 1234
DATA_SEED = 4321
STRIDE = 61


def batch_seed(step, rank=0):
    return DATA_SEED + 1000 * rank + step


def load(tag):
    return dict(np.load(os.path.join(GOLD, f"{tag}.npz"), allow_pickle=False))


def rel_l2(a, b):
    a = np.asarray(a, np.float64).ravel()
    b = np.asarray(b, np.float64).ravel()
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def check_fp(gold, prefix, name, arr, tol, what=""):
    """Compare a full tensor against a golden fingerprint: strided sample
    (relative L2), float64 sum and l2 norm."""
    arr = np.asarray(arr, np.float32).reshape(-1)
    smp = gold[f"{prefix}.{name}.sample"]
    assert smp.shape == arr[::STRIDE].shape, (what, prefix, name, smp.shape, arr[::STRIDE].shape)
    # error of the sampled elements relative to the tensor's RMS scale (a lone sampled element of a
    # tiny tensor can be far below the RMS, which would make a pure sample-relative figure noisy)
    scale = max(float(np.linalg.norm(smp.astype(np.float64))),
                float(gold[f"{prefix}.{name}.l2"]) * np.sqrt(smp.size / arr.size), 1e-30)
    e = float(np.linalg.norm(arr[::STRIDE].astype(np.float64) - smp.astype(np.float64))) / scale
    assert e <= tol, f"{what} {prefix}.{name}: sample relL2 {e:.3e} > {tol:.1e}"
    l2 = float(np.sqrt((arr.astype(np.float64) ** 2).sum()))
    assert abs(l2 - float(gold[f"{prefix}.{name}.l2"])) <= tol * max(l2, 1e-30) * 4, (what, prefix, name, "l2")
    s = float(arr.sum(dtype=np.float64))
    assert abs(s - float(gold[f"{prefix}.{name}.sum"])) <= tol * l2 * np.sqrt(arr.size) , (what, prefix, name, "sum")
