"""CPU oracle for the per-trial VAE training step (TEST INFRASTRUCTURE ONLY).

This file is a checker, not a product path: only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl
reference`` legs may import it.  Nothing under ``multidisttorch_b200/`` does.

What it restates (numpy, float32 arithmetic, explicit backward formulas --
no autograd):

* VAE forward                      ref:vae-hpo.py:29-45
* loss (BCE sum + KLD)             ref:vae-hpo.py:49-58
* backward of that loss            autograd of ref:vae-hpo.py:72
* DDP gradient averaging           ref:vae-hpo.py:130 (divide by n, then sum)
* Adam update                      ref:vae-hpo.py:131,74

The arithmetic itself lives in a third-party dependency that is absent from
/root/reference: PyTorch (no version pin in the reference; the oracle is pinned
against torch 2.11.0+cu128, the version installed in this image).  The
published semantics restated here:

* ``F.binary_cross_entropy``: each log is clamped at -100; backward is
  ``(r - x) / max((1 - r) * r, 1e-12)``.
* ``torch.sigmoid`` backward: ``g * (1 - y) * y``; relu backward passes
  gradient only where the *output* is > 0.
* ``torch.optim.Adam`` (no amsgrad / weight decay): ``m.lerp_(g, 1-b1)``,
  ``v = v*b2 + (1-b2)*g*g``, ``bc1 = 1-b1**t``, ``bc2 = 1-b2**t`` (Python
  doubles), ``denom = sqrt(v)/sqrt(bc2) + eps``, ``p -= (lr/bc1) * m/denom``.
* DDP default hook: ``bucket / n`` then all-reduce SUM.
* CUDA ``randn``: Philox4x32-10, ``curand_init(seed, subsequence=i, offset)``,
  lane .x of ``curand_normal4`` (Box-Muller).

Parity status: PINNED.  ``oracle/make_golden.py`` imports the real reference
(/root/reference/vae-hpo.py) in the build container, drives it with injected
seeds/eps, and writes ``tests/golden/*.npz``; ``tests/test_oracle_golden.py``
checks every function here against those vectors.

Flat parameter layout used by the B200 engine (and therefore by this oracle's
``pack``/``unpack``): W1[H,D] b1[H] W2x[2Z,H] b2x[2Z] W3[H,Z] b3[H] W4[D,H]
b4[D], where W2x = fc21.weight stacked over fc22.weight (mu rows first) and
b2x likewise.  Every segment starts on a 32-float boundary.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np

F32 = np.float32

SEGMENTS = ("W1", "b1", "W2x", "b2x", "W3", "b3", "W4", "b4")
ALIGN = 32  # floats


@dataclass(frozen=True)
class Dims:
    D: int = 784
    H: int = 400
    Z: int = 20

    def shapes(self):
        D, H, Z = self.D, self.H, self.Z
        return {
            "W1": (H, D), "b1": (H,),
            "W2x": (2 * Z, H), "b2x": (2 * Z,),
            "W3": (H, Z), "b3": (H,),
            "W4": (D, H), "b4": (D,),
        }

    def n_params(self):
        return sum(int(np.prod(s)) for s in self.shapes().values())


def _round_up(n, a):
    return (n + a - 1) // a * a


def layout(dims: Dims):
    """Segment name -> (offset, numel) in the flat buffer; plus total length."""
    off = 0
    out = {}
    for name in SEGMENTS:
        n = int(np.prod(dims.shapes()[name]))
        out[name] = (off, n)
        off += _round_up(n, ALIGN)
    return out, off


def pack(params: dict, dims: Dims) -> np.ndarray:
    lay, total = layout(dims)
    flat = np.zeros(total, dtype=F32)
    for name, (off, n) in lay.items():
        flat[off:off + n] = np.asarray(params[name], dtype=F32).reshape(-1)
    return flat


def unpack(flat: np.ndarray, dims: Dims) -> dict:
    lay, _ = layout(dims)
    shp = dims.shapes()
    return {k: flat[off:off + n].reshape(shp[k]) for k, (off, n) in lay.items()}


def init_params(seed: int, dims: Dims = Dims()) -> dict:
    """U(-1/sqrt(fan_in), +1/sqrt(fan_in)) for weight and bias, the same
    distribution nn.Linear uses (ref:vae-hpo.py:23-27).  Drawn from numpy's
    frozen legacy MT19937 stream so tests, bench and the golden generator get
    identical values without torch."""
    rs = np.random.RandomState(seed)
    D, H, Z = dims.D, dims.H, dims.Z

    def lin(out_f, in_f):
        k = 1.0 / math.sqrt(in_f)
        w = rs.uniform(-k, k, size=(out_f, in_f)).astype(F32)
        b = rs.uniform(-k, k, size=(out_f,)).astype(F32)
        return w, b

    W1, b1 = lin(H, D)
    W21, b21 = lin(Z, H)
    W22, b22 = lin(Z, H)
    W3, b3 = lin(H, Z)
    W4, b4 = lin(D, H)
    return {
        "W1": W1, "b1": b1,
        "W2x": np.concatenate([W21, W22], 0), "b2x": np.concatenate([b21, b22], 0),
        "W3": W3, "b3": b3, "W4": W4, "b4": b4,
    }


def make_batch(seed: int, B: int, dims: Dims = Dims()):
    """Synthetic MNIST-shaped batch x ~ U[0,1) and eps ~ N(0,1), float32."""
    rs = np.random.RandomState(seed)
    x = rs.random_sample((B, dims.D)).astype(F32)
    eps = rs.standard_normal((B, dims.Z)).astype(F32)
    return x, eps


# ----------------------------------------------------------------------------
# forward / loss / backward
# ----------------------------------------------------------------------------

def sigmoid(a):
    a = a.astype(F32)
    return (F32(1) / (F32(1) + np.exp(-a, dtype=F32))).astype(F32)


def forward(p: dict, x: np.ndarray, eps: np.ndarray, dims: Dims = Dims()) -> dict:
    """ref:vae-hpo.py:29-45.  x is [B, D] (the caller flattens 1x28x28)."""
    Z = dims.Z
    x = np.ascontiguousarray(x, dtype=F32).reshape(-1, dims.D)
    a1 = x @ p["W1"].T + p["b1"]
    h1 = np.maximum(a1, F32(0))
    ml = h1 @ p["W2x"].T + p["b2x"]
    mu, lv = ml[:, :Z], ml[:, Z:]
    std = np.exp(F32(0.5) * lv, dtype=F32)
    z = mu + eps.astype(F32) * std
    a3 = z @ p["W3"].T + p["b3"]
    h3 = np.maximum(a3, F32(0))
    a4 = h3 @ p["W4"].T + p["b4"]
    r = sigmoid(a4)
    return dict(x=x, h1=h1, mu=mu, lv=lv, std=std, eps=eps.astype(F32), z=z, h3=h3, a4=a4, recon=r)


def loss_terms(recon, x, mu, lv):
    """ref:vae-hpo.py:49-58 -> (BCE, KLD), each a float64 accumulation of
    float32 element terms (the reference's reduction order is unspecified)."""
    r = recon.astype(F32)
    x = x.astype(F32)
    with np.errstate(divide="ignore"):
        log_r = np.maximum(np.log(r, dtype=F32), F32(-100))
        log_1mr = np.maximum(np.log1p(-r, dtype=F32), F32(-100))
    bce_el = (x - F32(1)) * log_1mr - x * log_r
    kld_el = F32(1) + lv - mu * mu - np.exp(lv, dtype=F32)
    return float(bce_el.sum(dtype=np.float64)), float(-0.5 * kld_el.sum(dtype=np.float64))


def loss_function(recon, x, mu, lv) -> float:
    b, k = loss_terms(recon, x, mu, lv)
    return b + k


def loss_grads(recon, x, mu, lv):
    """d loss / d(recon, mu, logvar): torch's binary_cross_entropy_backward
    plus the closed-form KLD derivative."""
    r = recon.astype(F32)
    x = x.astype(F32)
    d_recon = (r - x) / np.maximum((F32(1) - r) * r, F32(1e-12))
    d_mu = mu.astype(F32).copy()
    d_lv = F32(0.5) * (np.exp(lv, dtype=F32) - F32(1))
    return d_recon.astype(F32), d_mu, d_lv.astype(F32)


def backward_from(p: dict, act: dict, d_recon, d_mu, d_lv, dims: Dims = Dims()) -> dict:
    """Back-propagate arbitrary upstream gradients of (recon, mu, logvar)
    through decode / reparameterize / encode.  SURVEY App. A-2."""
    Z = dims.Z
    r, h3, z, h1, x = act["recon"], act["h3"], act["z"], act["h1"], act["x"]
    da4 = (d_recon * (F32(1) - r) * r).astype(F32)          # sigmoid_backward
    g = {}
    g["W4"] = da4.T @ h3
    g["b4"] = da4.sum(0, dtype=F32)
    dh3 = da4 @ p["W4"]
    da3 = np.where(h3 > 0, dh3, F32(0)).astype(F32)          # threshold_backward
    g["W3"] = da3.T @ z
    g["b3"] = da3.sum(0, dtype=F32)
    dz = da3 @ p["W3"]
    dmu = dz + d_mu
    dlv = dz * act["eps"] * act["std"] * F32(0.5) + d_lv
    dml = np.concatenate([dmu, dlv], 1).astype(F32)
    g["W2x"] = dml.T @ h1
    g["b2x"] = dml.sum(0, dtype=F32)
    dh1 = dml @ p["W2x"]
    da1 = np.where(h1 > 0, dh1, F32(0)).astype(F32)
    g["W1"] = da1.T @ x
    g["b1"] = da1.sum(0, dtype=F32)
    return {k: np.ascontiguousarray(v, dtype=F32) for k, v in g.items()}


def loss_and_grads(p: dict, x, eps, dims: Dims = Dims()):
    act = forward(p, x, eps, dims)
    L = loss_function(act["recon"], act["x"], act["mu"], act["lv"])
    d_recon, d_mu, d_lv = loss_grads(act["recon"], act["x"], act["mu"], act["lv"])
    g = backward_from(p, act, d_recon, d_mu, d_lv, dims)
    return L, g, act


# ----------------------------------------------------------------------------
# DDP averaging + Adam
# ----------------------------------------------------------------------------

def ddp_average(grads_per_rank: list) -> dict:
    """Default DDP comm hook: every rank divides by n, then SUM all-reduce
    (accumulated here in rank order)."""
    n = len(grads_per_rank)
    out = {}
    for k in grads_per_rank[0]:
        acc = grads_per_rank[0][k] / F32(n)
        for r in range(1, n):
            acc = acc + grads_per_rank[r][k] / F32(n)
        out[k] = acc.astype(F32)
    return out


def adam_scalars(t: int, lr=1e-3, b1=0.9, b2=0.999):
    """The scalars torch computes in Python doubles before touching tensors."""
    bc1 = 1.0 - b1 ** t
    bc2 = 1.0 - b2 ** t
    return lr / bc1, math.sqrt(bc2)


def adam_update(p, g, m, v, t, lr=1e-3, b1=0.9, b2=0.999, eps=1e-8):
    """In-place single-tensor Adam step t (1-based) on float32 arrays."""
    step_size, bc2_sqrt = adam_scalars(t, lr, b1, b2)
    g = g.astype(F32)
    m += (g - m) * F32(1.0 - b1)
    v *= F32(b2)
    v += F32(1.0 - b2) * g * g
    denom = np.sqrt(v, dtype=F32) / F32(bc2_sqrt) + F32(eps)
    p -= F32(step_size) * (m / denom)


def sgd_update(p, g, buf, t, lr, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False):
    """In-place torch.optim.SGD step t (1-based) on float32 arrays (torch/optim/sgd.py _single_tensor_sgd):
    g += wd*p; buf = g on the first step, else momentum*buf + (1-dampening)*g; g = g + momentum*buf (nesterov)
    or buf; p -= lr*g.  BASELINE north_star (c) names SGD next to Adam; the reference itself uses Adam only."""
    g = g.astype(F32)
    if weight_decay != 0:
        g = g + F32(weight_decay) * p
    if momentum != 0:
        if t == 1:
            buf[...] = g
        else:
            buf *= F32(momentum)
            buf += F32(1.0 - dampening) * g
        g = g + F32(momentum) * buf if nesterov else buf
    p -= F32(lr) * g


class Trainer:
    """N-step driver mirroring ref:vae-hpo.py:69-74 for one rank (or, with
    several batches per step, one DDP group)."""

    def __init__(self, params: dict, dims: Dims = Dims(), lr=1e-3, b1=0.9, b2=0.999, eps=1e-8, sgd=None):
        self.sgd = sgd          # dict(lr=, momentum=, dampening=, weight_decay=, nesterov=) -> SGD instead of Adam
        self.dims = dims
        self.p = {k: np.array(v, dtype=F32, copy=True) for k, v in params.items()}
        self.m = {k: np.zeros_like(v) for k, v in self.p.items()}
        self.v = {k: np.zeros_like(v) for k, v in self.p.items()}
        self.t = 0
        self.lr, self.b1, self.b2, self.eps = lr, b1, b2, eps

    def step(self, batches):
        """batches: list of (x, eps), one per rank of the group.  Returns the
        per-rank losses and the averaged gradients."""
        losses, grads = [], []
        for x, eps in batches:
            L, g, _ = loss_and_grads(self.p, x, eps, self.dims)
            losses.append(L)
            grads.append(g)
        g = grads[0] if len(grads) == 1 else ddp_average(grads)
        self.t += 1
        for k in self.p:
            if self.sgd is not None:
                sgd_update(self.p[k], g[k], self.m[k], self.t, **self.sgd)
            else:
                adam_update(self.p[k], g[k], self.m[k], self.v[k], self.t, lr=self.lr, b1=self.b1, b2=self.b2, eps=self.eps)
        return losses, g


# ----------------------------------------------------------------------------
# Philox4x32-10 + Box-Muller, as curand / torch-CUDA randn use them
# ----------------------------------------------------------------------------

_PH_M0 = np.uint64(0xD2511F53)
_PH_M1 = np.uint64(0xCD9E8D57)
_PH_W0 = np.uint32(0x9E3779B9)
_PH_W1 = np.uint32(0xBB67AE85)


def philox4x32_10(ctr: np.ndarray, key: np.ndarray) -> np.ndarray:
    """ctr: [n,4] uint32, key: [2] uint32 -> [n,4] uint32."""
    c = ctr.astype(np.uint32).copy()
    k0 = np.uint32(key[0])
    k1 = np.uint32(key[1])
    with np.errstate(over="ignore"):
        for _ in range(10):
            p0 = c[:, 0].astype(np.uint64) * _PH_M0
            p1 = c[:, 2].astype(np.uint64) * _PH_M1
            hi0 = (p0 >> np.uint64(32)).astype(np.uint32)
            lo0 = p0.astype(np.uint32)
            hi1 = (p1 >> np.uint64(32)).astype(np.uint32)
            lo1 = p1.astype(np.uint32)
            n0 = hi1 ^ c[:, 1] ^ k0
            n1 = lo1
            n2 = hi0 ^ c[:, 3] ^ k1
            n3 = lo0
            c = np.stack([n0, n1, n2, n3], 1)
            k0 = np.uint32(k0 + _PH_W0)
            k1 = np.uint32(k1 + _PH_W1)
    return c


def cuda_randn(seed: int, offset: int, numel: int) -> np.ndarray:
    """Values torch.randn(numel, device='cuda') yields for a generator at
    (seed, offset) when numel fits one grid pass (numel <= 256*blocks): element
    i = lane .x of curand_normal4 on Philox(seed, subsequence=i, offset).
    float32 math follows curand_normal.h (_curand_box_muller)."""
    assert offset % 4 == 0
    i = np.arange(numel, dtype=np.uint64)
    n = np.uint64(offset // 4)
    ctr = np.zeros((numel, 4), dtype=np.uint32)
    ctr[:, 0] = np.uint32(n & np.uint64(0xFFFFFFFF))
    ctr[:, 1] = np.uint32(n >> np.uint64(32))
    ctr[:, 2] = (i & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    ctr[:, 3] = (i >> np.uint64(32)).astype(np.uint32)
    key = np.array([seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF], dtype=np.uint32)
    r = philox4x32_10(ctr, key)
    two_pow32_inv = F32(2.3283064e-10)
    two_pow32_inv_2pi = F32(2.3283064e-10 * 6.2831855)
    u = r[:, 0].astype(F32) * two_pow32_inv + F32(two_pow32_inv / F32(2))
    v = r[:, 1].astype(F32) * two_pow32_inv_2pi + F32(two_pow32_inv_2pi / F32(2))
    s = np.sqrt(F32(-2) * np.log(u, dtype=F32), dtype=F32)
    return (s * np.sin(v, dtype=F32)).astype(F32)


# ----------------------------------------------------------------------------
# fingerprints for compact golden files
# ----------------------------------------------------------------------------

def fingerprint(a: np.ndarray, stride: int = 61) -> dict:
    a = np.asarray(a, dtype=F32).reshape(-1)
    return {
        "sum": np.float64(a.sum(dtype=np.float64)),
        "l2": np.float64(np.sqrt((a.astype(np.float64) ** 2).sum())),
        "sample": a[::stride].copy(),
    }
