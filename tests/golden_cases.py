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
    (relative L2 against the tensor's RMS scale), float64 sum and l2 norm."""
    from multidisttorch_b200.synth import fingerprint_error
    e, e_l2, e_sum = fingerprint_error(gold, prefix, name, arr)
    assert e <= tol, f"{what} {prefix}.{name}: sample relL2 {e:.3e} > {tol:.1e}"
    assert e_l2 <= 4 * tol, (what, prefix, name, "l2", e_l2)
    assert e_sum <= tol, (what, prefix, name, "sum", e_sum)
    return e
