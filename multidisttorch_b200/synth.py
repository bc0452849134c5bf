"""Seeded synthetic inputs of the VAE-HPO step: initial parameters and MNIST-shaped batches.

bench.py, the drop-in self-check and the golden replay (`golden.py`) need the same
deterministic parameters / batches the test fixtures were generated with, without importing
test infrastructure.  Everything here is input generation from numpy's frozen legacy MT19937
stream -- no model arithmetic.  tests/test_synth.py pins these generators bit-for-bit against
the ones the golden fixtures were produced with.
"""
import math

import numpy as np

SEGMENTS = ("W1", "b1", "W2x", "b2x", "W3", "b3", "W4", "b4")
PARAM_SEED = 1234
DATA_SEED = 4321


def init_params(seed, D=784, H=400, Z=20):
    """U(-1/sqrt(fan_in), 1/sqrt(fan_in)) for weights and biases, the distribution nn.Linear draws from
    (ref:vae-hpo.py:23-27), in the engine's segment naming (W2x = fc21 stacked over fc22)."""
    rs = np.random.RandomState(seed)

    def lin(out_f, in_f):
        k = 1.0 / math.sqrt(in_f)
        w = rs.uniform(-k, k, size=(out_f, in_f)).astype(np.float32)
        b = rs.uniform(-k, k, size=(out_f,)).astype(np.float32)
        return w, b

    W1, b1 = lin(H, D)
    W21, b21 = lin(Z, H)
    W22, b22 = lin(Z, H)
    W3, b3 = lin(H, Z)
    W4, b4 = lin(D, H)
    return {"W1": W1, "b1": b1, "W2x": np.concatenate([W21, W22], 0), "b2x": np.concatenate([b21, b22], 0),
            "W3": W3, "b3": b3, "W4": W4, "b4": b4}


def make_batch(seed, B, D=784, Z=20):
    """x ~ U[0,1) of shape [B, D] and eps ~ N(0,1) of shape [B, Z], float32."""
    rs = np.random.RandomState(seed)
    x = rs.random_sample((B, D)).astype(np.float32)
    eps = rs.standard_normal((B, Z)).astype(np.float32)
    return x, eps


def batch_seed(step, rank=0):
    """Seed of the batch world-rank `rank` sees at step `step` in the golden fixtures."""
    return DATA_SEED + 1000 * rank + step


# ---------------------------------------------------------------------------- golden fingerprints
def fingerprint_error(gold, prefix, name, arr):
    """Error of tensor `arr` against the fingerprint a golden file holds for `prefix.name`: every `stride`-th
    element (relative L2 against the tensor's RMS scale), the float64 sum and the l2 norm.
    Returns (sample_rel_l2, l2_rel_err, sum_err_over_l2_sqrt_n)."""
    arr = np.asarray(arr, np.float32).reshape(-1)
    stride = int(gold["stride"]) if "stride" in gold else 61
    smp = gold[f"{prefix}.{name}.sample"]
    got = arr[::stride]
    if smp.shape != got.shape:
        raise ValueError(f"{prefix}.{name}: fingerprint has {smp.shape} samples, tensor gives {got.shape}")
    g_l2 = float(gold[f"{prefix}.{name}.l2"])
    scale = max(float(np.linalg.norm(smp.astype(np.float64))), g_l2 * np.sqrt(smp.size / arr.size), 1e-30)
    e = float(np.linalg.norm(got.astype(np.float64) - smp.astype(np.float64))) / scale
    l2 = float(np.sqrt((arr.astype(np.float64) ** 2).sum()))
    e_l2 = abs(l2 - g_l2) / max(l2, 1e-30)
    s = float(arr.sum(dtype=np.float64))
    e_sum = abs(s - float(gold[f"{prefix}.{name}.sum"])) / max(l2 * np.sqrt(arr.size), 1e-30)
    return e, e_l2, e_sum
