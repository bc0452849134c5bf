"""GPU parity tests: every call goes through the C-ABI (ctypes) into the sm_100a
kernels and is compared with the CPU oracle and with the golden vectors the real
reference produced.  Run with `pytest -m gpu` on the B200 box."""
import os

import numpy as np
import pytest
import torch

from oracle import vae_numpy as O
from tests.golden_cases import PARAM_SEED, batch_seed, check_fp, load, rel_l2

pytestmark = pytest.mark.gpu

# Stated tolerances (relative L2 against the reference's golden fingerprints; loss relative).  Every bound is at most
# ~3x the largest error MEASURED on a B200 over the five golden training cases (B = 128 / 48 / 1 / 512 at H = 400 and
# B = 128 at H = 4096; scripts/measure_errors.py -> profiles/r2_measured_errors.txt), the measured maxima in brackets:
#   fp32   = true fp32 FFMA GEMMs: differs from the reference (torch fp32 on CPU) by summation order and FMA contraction
#            [loss 2.2e-7, grad 7.6e-7, param 1.5e-7, m 6.7e-7, v 1.2e-6]
#   tf32x3 = 3-pass split TF32 on tcgen05, fp32-grade  [loss 4.9e-7, grad 2.9e-6, param 3.5e-7, m 3.6e-6, v 6.9e-6]
#   tf32   = single pass, 10-bit operand mantissas: loss-level and single-step-gradient-level agreement only
#            [loss 5.1e-4, grad 9.5e-3, param 9.8e-3, m 4.1e-2, v 1.7e-2]
#   bf16   = bf16 operands (8-bit mantissa), fp32 accumulate; Adam's normalisation turns the gradient noise into O(lr)
#            parameter differences, so parameters / moments are checked loosely (SURVEY section 7 "Precision vs parity")
#            [loss 1.7e-4, grad 2.3e-2, param 1.8e-2, m 1.5e-1, v 6.4e-2]
# `act` bounds single GEMM outputs (test_linear_vs_oracle, forward goldens); `mom` bounds exp_avg, 2 x mom exp_avg_sq.
TOL = {
    "fp32":   dict(loss=1e-6, act=2e-6, grad=3e-6, param=1e-6, mom=2e-6),
    "tf32x3": dict(loss=2e-6, act=2e-5, grad=1e-5, param=2e-6, mom=1.2e-5),
    "tf32":   dict(loss=2e-3, act=5e-3, grad=3e-2, param=3e-2, mom=1.2e-1),
    "bf16":   dict(loss=6e-4, act=2e-2, grad=7e-2, param=6e-2, mom=4.5e-1),
}
MODES = [m for m in os.environ.get("MDT_TEST_MODES", "fp32,tf32x3,tf32,bf16").split(",") if m]
LINEAR_MODES = [m for m in MODES if m != "bf16"]       # mdt_linear takes fp32 operands; bf16 lives inside the engine


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.fixture(scope="module")
def E():
    from multidisttorch_b200 import engine
    assert torch.cuda.is_available()
    return engine


def make_engine(E, mode, B, H=400, **kw):
    eng = E.VAEEngine(hidden=H, max_batch=max(B, 1), gemm_mode=mode, **kw)
    eng.load_segments(O.init_params(PARAM_SEED, O.Dims(H=H)))
    return eng


# ------------------------------------------------------------------ stand-alone kernels

def test_adam_flat_matches_oracle(E):
    rs = np.random.RandomState(0)
    n = 652896
    p = rs.standard_normal(n).astype(np.float32) * 0.05
    m = np.zeros(n, np.float32)
    v = np.zeros(n, np.float32)
    dp, dm, dv = dev(p), dev(m), dev(v)
    for t in range(1, 6):
        g = (rs.standard_normal(n) * 10.0 ** rs.uniform(-6, 1, n)).astype(np.float32)
        O.adam_update(p, g * np.float32(0.5), m, v, t)
        E.adam_flat(dp, dev(g), dm, dv, t, grad_scale=0.5)
    torch.cuda.synchronize()
    assert rel_l2(dm.cpu().numpy(), m) < 1e-6
    assert rel_l2(dv.cpu().numpy(), v) < 1e-6
    assert rel_l2(dp.cpu().numpy(), p) < 1e-6
    assert np.abs(dp.cpu().numpy() - p).max() < 1e-6


def test_adam_flat_odd_sizes_and_errors(E):
    for n in (4, 8, 1028):
        p, g, m, v = (torch.zeros(n, device="cuda") for _ in range(4))
        g.fill_(1.0)
        E.adam_flat(p, g, m, v, 1)
        torch.cuda.synchronize()
        np.testing.assert_allclose(p.cpu().numpy(), -1e-3, rtol=1e-5)
    with pytest.raises(RuntimeError, match="1-based"):
        E.adam_flat(p, g, m, v, 0)


def test_loss_kernel_golden_corners(E):
    g = load("loss_edge")
    loss, dr, dm, dl = E.loss_bce_kl(dev(g["recon"]), dev(g["x"]), dev(g["mu"]), dev(g["logvar"]))
    torch.cuda.synchronize()
    assert abs(float(loss.item()) - float(g["loss"])) <= 2e-6 * abs(float(g["loss"]))
    np.testing.assert_allclose(dr.cpu().numpy(), g["d_recon"], rtol=3e-6, atol=0)
    np.testing.assert_array_equal(dm.cpu().numpy(), g["d_mu"])
    np.testing.assert_allclose(dl.cpu().numpy(), g["d_logvar"], rtol=3e-6, atol=1.2e-7)


@pytest.mark.parametrize("B", [1, 3, 128, 1000])
def test_loss_kernel_random_vs_oracle(E, B):
    rs = np.random.RandomState(B)
    recon = rs.random_sample((B, 784)).astype(np.float32)
    x = rs.random_sample((B, 784)).astype(np.float32)
    mu = rs.standard_normal((B, 20)).astype(np.float32)
    lv = rs.standard_normal((B, 20)).astype(np.float32)
    loss, dr, dm, dl = E.loss_bce_kl(dev(recon), dev(x), dev(mu), dev(lv))
    loss2, _, _, _ = E.loss_bce_kl(dev(recon), dev(x), dev(mu), dev(lv), want_grads=False)
    ref = O.loss_function(recon, x, mu, lv)
    rdr, rdm, rdl = O.loss_grads(recon, x, mu, lv)
    assert abs(float(loss.item()) - ref) <= 2e-6 * abs(ref)
    assert float(loss.item()) == float(loss2.item())          # fixed summation order: bit-reproducible
    assert rel_l2(dr.cpu().numpy(), rdr) < 1e-6 and rel_l2(dl.cpu().numpy(), rdl) < 1e-6
    np.testing.assert_array_equal(dm.cpu().numpy(), rdm)


def test_reparam_supplied_eps(E):
    rs = np.random.RandomState(1)
    ml = rs.standard_normal((128, 40)).astype(np.float32)
    eps = rs.standard_normal((128, 20)).astype(np.float32)
    dml = dev(ml)
    z, eo = E.reparam(dml[:, :20], dml[:, 20:], dev(eps))
    ref = ml[:, :20] + eps * np.exp(np.float32(0.5) * ml[:, 20:], dtype=np.float32)
    np.testing.assert_allclose(z.cpu().numpy(), ref, rtol=2e-6, atol=5e-7)  # mu + eps*std cancels near 0
    np.testing.assert_array_equal(eo.cpu().numpy(), eps)


@pytest.mark.parametrize("B,offset", [(128, 0), (128, 4 * 37), (1, 8), (20000, 0)])
def test_reparam_philox_is_torch_cuda_randn(E, B, offset):
    """eps drawn inside the kernel must be bit-identical to what the reference's
    torch.randn_like(std) yields on CUDA for the same generator state (ref:vae-hpo.py:35)."""
    seed = 1234
    gen = torch.Generator(device="cuda")
    gen.manual_seed(seed)
    gen.set_offset(offset)
    want = torch.randn(B, 20, device="cuda", generator=gen)
    mu = torch.zeros(B, 20, device="cuda")
    lv = torch.zeros(B, 20, device="cuda")
    z, eps = E.reparam(mu, lv, None, seed=seed, offset=offset)
    torch.cuda.synchronize()
    assert torch.equal(eps, want)
    assert torch.equal(z, want)           # mu=0, std=1
    if B <= 128:                          # and the numpy restatement of Philox/Box-Muller agrees (libm sin vs __sincosf)
        np.testing.assert_allclose(O.cuda_randn(seed, offset, B * 20), want.cpu().numpy().ravel(), rtol=0, atol=2e-5)


@pytest.mark.parametrize("mode", LINEAR_MODES)
@pytest.mark.parametrize("M,N,K,act", [(128, 400, 784, "relu"), (128, 40, 400, "none"), (128, 400, 20, "relu"),
                                       (128, 784, 400, "sigmoid"), (48, 400, 784, "relu"), (1, 784, 400, "none"),
                                       (300, 400, 784, "none")])
def test_linear_vs_oracle(E, mode, M, N, K, act):
    rs = np.random.RandomState(M + N + K)
    x = rs.random_sample((M, K)).astype(np.float32)
    w = (rs.standard_normal((N, K)) / np.sqrt(K)).astype(np.float32)
    b = rs.standard_normal(N).astype(np.float32)
    y = E.linear(dev(x), dev(w), dev(b), act=act, gemm_mode=mode).cpu().numpy()
    ref = x.astype(np.float64) @ w.astype(np.float64).T + b
    ref = {"none": ref, "relu": np.maximum(ref, 0), "sigmoid": 1 / (1 + np.exp(-ref))}[act]
    assert rel_l2(y, ref) < TOL[mode]["act"], rel_l2(y, ref)


# ------------------------------------------------------------------ engine vs golden

@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("tag", ["step_b128", "step_b48", "step_b1", "step_b512", "step_h4096_b128"])
@pytest.mark.parametrize("graph", [True, False])
def test_train_steps_match_reference_golden(E, mode, tag, graph):
    """step_b512 = BASELINE config 4's group batch on one rank; step_h4096_b128 = config 5's width (the reference's
    VAE with its five nn.Linear re-sized, oracle/make_golden.py:widen)."""
    g = load(tag)
    tol = TOL[mode]
    B, n = int(g["B"]), int(g["n_steps"])
    H = int(g["H"]) if "H" in g else 400
    if H != 400 and (not graph or mode == "fp32"):
        pytest.skip("the wide case runs once per tensor-core mode (graph on); fp32 FFMA at H=4096 is covered by test_hidden_4096")
    eng = make_engine(E, mode, B, H=H, use_graph=graph)
    for t in range(n):
        x, eps = O.make_batch(batch_seed(t), B, O.Dims(H=H))
        loss = eng.train_step(dev(x), dev(eps))
        got = float(loss.item())
        assert abs(got - g["losses"][t]) <= tol["loss"] * abs(g["losses"][t]), (tag, t, got, g["losses"][t])
        if t == 0:
            gr = eng.segments_to_numpy(eng.grads)
            for k in O.SEGMENTS:
                check_fp(g, "grad0", k, gr[k], tol["grad"], f"{tag}/{mode}")
    pN = eng.segments_to_numpy()
    mN = eng.segments_to_numpy(eng.exp_avg)
    vN = eng.segments_to_numpy(eng.exp_avg_sq)
    for k in O.SEGMENTS:
        check_fp(g, "paramN", k, pN[k], tol["param"], f"{tag}/{mode}")
        check_fp(g, "mN", k, mN[k], tol["mom"], f"{tag}/{mode}")
        check_fp(g, "vN", k, vN[k], 2 * tol["mom"], f"{tag}/{mode}")
    assert eng.state()[0] == n
    eng.close()


@pytest.mark.parametrize("mode", MODES)
def test_fused_bce_epilogue_on_reference_corners(E, mode):
    """The training step computes sigmoid + BCE + d(logits) inside fc4's GEMM epilogue with SFU approximations
    (gemm_tcgen05.cuh fast_sigmoid / fast_bce_term / fast_dlogit), not with the stand-alone loss kernel.  Here fc4's
    logits are pinned to the clamp / saturation corners of F.binary_cross_entropy and torch.sigmoid (|a| up to 120:
    r = 0, denormal r, r = 1 - 2^-24, r = 1) and the step is compared with what the reference produced there
    (oracle/make_golden.py:run_step_edge)."""
    from oracle.make_golden import EDGE_LOGITS, edge_setup
    g = load("step_edge_b8")
    tol = TOL[mode]
    B, n = int(g["B"]), len(EDGE_LOGITS)
    x, eps = O.make_batch(batch_seed(0), B)
    p0, x = edge_setup(O.init_params(PARAM_SEED), x)
    eng = E.VAEEngine(max_batch=B, gemm_mode=mode)
    eng.load_segments(p0)
    loss = float(eng.train_step(dev(x), dev(eps)).item())
    assert abs(loss - g["losses"][0]) <= tol["loss"] * abs(g["losses"][0]), (loss, g["losses"][0])
    recon = eng.workspace("recon", B).cpu().numpy()[:3, :n]
    # recon at the corners: exact zeros and ones where the reference has them; elsewhere (denormals included) within the
    # SFU forms' error -- ex2.approx and the range reduction by log2(e) are ~2 ulp each: measured 2.4e-6 relative
    want = g["recon_edge"]
    assert np.array_equal(recon == 0, want == 0) and np.array_equal(recon == 1, want == 1), (recon, want)
    np.testing.assert_allclose(recon, want, rtol=6e-6, atol=1e-44)
    gr = eng.segments_to_numpy(eng.grads)
    # d loss / d fc4.bias at the pinned logits = sum over the batch of d(logits): exercises fast_dlogit's clamp branch
    # (bf16 mode sums the bias gradient from the bf16 operand copy of d(logits): 2^-9 per element)
    np.testing.assert_allclose(gr["b4"][:n], g["grad0.b4.edge"], rtol=2e-5 if mode != "bf16" else 5e-3, atol=1e-30)
    for k in O.SEGMENTS:
        check_fp(g, "grad0", k, gr[k], tol["grad"], f"edge/{mode}")
    pN = eng.segments_to_numpy()
    for k in O.SEGMENTS:
        check_fp(g, "paramN", k, pN[k], tol["param"], f"edge/{mode}")
    eng.close()


@pytest.mark.parametrize("mode", [m for m in MODES if m != "fp32"])
@pytest.mark.parametrize("env", [{"MDT_TC_LATENT": "1"}, {"MDT_NO_FUSED_ADAM": "1"},
                                 {"MDT_TC_LATENT": "1", "MDT_NO_FUSED_ADAM": "1"}, {"MDT_LATENT_SC": "1"},
                                 {"MDT_LATENT_SC": "2"}, {"MDT_DW4_LATE": "1"}, {"MDT_W4_SPLIT": "1"}, {"MDT_A_TMEM": "1"},
                                 {"MDT_NO_CLUSTER_SPLITK": "1"}],
                         ids=lambda e: "+".join(f"{k[4:].lower()}{v}" for k, v in e.items()))
def test_alternative_kernel_paths_match_golden(E, mode, env, monkeypatch):
    """The step has run-time alternatives the defaults do not take at this size (tensor-core latent layers, the
    flat Adam kernel, other samples-per-CTA / launch orders): each must meet the same golden vectors."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    g = load("step_b48")
    tol = TOL[mode]
    B, n = int(g["B"]), int(g["n_steps"])
    eng = make_engine(E, mode, B)
    for t in range(n):
        x, eps = O.make_batch(batch_seed(t), B)
        loss = float(eng.train_step(dev(x), dev(eps)).item())
        assert abs(loss - g["losses"][t]) <= tol["loss"] * abs(g["losses"][t]), (env, t, loss, g["losses"][t])
        if t == 0:
            gr = eng.segments_to_numpy(eng.grads)
            for k in O.SEGMENTS:
                check_fp(g, "grad0", k, gr[k], tol["grad"], f"{env}/{mode}")
    pN = eng.segments_to_numpy()
    for k in O.SEGMENTS:
        check_fp(g, "paramN", k, pN[k], tol["param"], f"{env}/{mode}")
    assert eng.state()[0] == n
    eng.close()


@pytest.mark.parametrize("mode", MODES)
def test_forward_eval_decode_golden(E, mode):
    g = load("eval_b64")
    tol = TOL[mode]
    B = int(g["B"])
    eng = make_engine(E, mode, B)
    x, eps = O.make_batch(batch_seed(77), B)
    recon, mu, lv = eng.forward(dev(x), dev(eps))
    check_fp(g, "eval", "recon", recon.cpu().numpy(), tol["act"])
    check_fp(g, "eval", "mu", mu.cpu().numpy(), tol["act"])
    check_fp(g, "eval", "logvar", lv.cpu().numpy(), tol["act"])
    loss, rec2 = eng.eval_step(dev(x), dev(eps), want_recon=True)
    assert abs(float(loss.item()) - float(g["loss"])) <= tol["loss"] * abs(float(g["loss"]))
    assert torch.equal(rec2, recon)
    zs = np.random.RandomState(99).standard_normal((64, 20)).astype(np.float32)
    check_fp(g, "eval", "decode64", eng.decode(dev(zs)).cpu().numpy(), tol["act"])
    eng.close()


@pytest.mark.parametrize("mode", MODES)
def test_compat_mode_backward_from_torch_loss(E, mode):
    """Compat mode: the script's own loss_function runs in torch on the
    engine's outputs; mdt_backward consumes dL/d(recon, mu, logvar)."""
    tol = TOL[mode]
    B = 64
    eng = make_engine(E, mode, B)
    p0 = O.init_params(PARAM_SEED)
    x, eps = O.make_batch(5, B)
    recon, mu, lv = eng.forward(dev(x), dev(eps))
    recon.requires_grad_(); mu.requires_grad_(); lv.requires_grad_()
    bce = torch.nn.functional.binary_cross_entropy(recon, dev(x), reduction="sum")
    kld = -0.5 * torch.sum(1 + lv - mu.pow(2) - lv.exp())
    (bce + kld).backward()
    eng.backward(recon.grad, mu.grad, lv.grad)
    L, gref, _ = O.loss_and_grads(p0, x, eps)
    assert abs(float((bce + kld).item()) - L) <= tol["loss"] * abs(L)
    got = eng.segments_to_numpy(eng.grads)
    for k in O.SEGMENTS:
        assert rel_l2(got[k], gref[k]) < tol["grad"], (k, rel_l2(got[k], gref[k]))
    # then allreduce (no-op for a 1-rank group) + Adam, like DDP + optimizer.step()
    eng.allreduce_grads()
    eng.adam_step()
    tr = O.Trainer(p0)
    tr.step([(x, eps)])
    pN = eng.segments_to_numpy()
    for k in O.SEGMENTS:
        assert rel_l2(pN[k], tr.p[k]) < tol["param"], k
    eng.close()


@pytest.mark.parametrize("mode", MODES)
def test_host_buffer_step_equals_device_step(E, mode):
    B = 128
    a = make_engine(E, mode, B)
    b = make_engine(E, mode, B)
    for t in range(3):
        x, eps = O.make_batch(batch_seed(t), B)
        la = float(a.train_step(dev(x), dev(eps)).item())
        lb = b.train_step_host(torch.from_numpy(x).pin_memory(), torch.from_numpy(eps).pin_memory())
        assert la == lb
    assert torch.equal(a.params, b.params)
    a.close(); b.close()


@pytest.mark.parametrize("mode", MODES)
def test_philox_step_uses_and_advances_device_state(E, mode):
    B = 128
    eng = make_engine(E, mode, B)
    eng.set_rng(99, 40)
    x, _ = O.make_batch(1, B)
    loss = eng.train_step(dev(x), None)
    used = eng.workspace("eps", B)
    gen = torch.Generator(device="cuda"); gen.manual_seed(99); gen.set_offset(40)
    assert torch.equal(used, torch.randn(B, 20, device="cuda", generator=gen))
    assert eng.state() == (1, 44)
    # the same step with that eps supplied explicitly gives the same loss on a fresh engine
    eng2 = make_engine(E, mode, B)
    assert float(eng2.train_step(dev(x), used).item()) == float(loss.item())
    eng.train_step(dev(x), None)
    assert eng.state() == (2, 48)
    eng.close(); eng2.close()


def test_empty_batch_and_errors(E):
    eng = make_engine(E, "fp32", 16)
    before = eng.params.clone()
    loss = eng.train_step(torch.empty(0, 784, device="cuda"), None)
    assert float(loss.item()) == 0.0 and torch.equal(before, eng.params) and eng.state()[0] == 0
    with pytest.raises(RuntimeError, match="max_batch"):
        eng.train_step(torch.zeros(17, 784, device="cuda"), None)
    eng.close()


@pytest.mark.parametrize("mode", MODES)
def test_full_size_properties(E, mode):
    """At BASELINE size (B=128, 784-400-20), size-independent properties: bit-reproducible runs,
    graph == eager, loss goes down on a fixed batch, step counter and Philox offset advance."""
    B = 128
    x, _ = O.make_batch(3, B)
    runs = []
    for graph in (True, True, False):
        eng = make_engine(E, mode, B, use_graph=graph)
        eng.set_rng(5, 0)
        losses = [float(eng.train_step(dev(x), None).item()) for _ in range(40)]
        runs.append((losses, eng.params.clone()))
        assert eng.state() == (40, 160)
        eng.close()
    assert runs[0][0] == runs[1][0] and torch.equal(runs[0][1], runs[1][1])
    assert runs[0][0] == runs[2][0] and torch.equal(runs[0][1], runs[2][1])
    assert runs[0][0][-1] < 0.99 * runs[0][0][0]
    assert np.isfinite(runs[0][0]).all()


@pytest.mark.parametrize("mode", MODES)
def test_hidden_4096_one_step_vs_oracle(E, mode):
    """BASELINE config 5 shape (H=4096), one step against the oracle."""
    dims = O.Dims(H=4096)
    B = 32
    p0 = O.init_params(11, dims)
    eng = E.VAEEngine(hidden=4096, max_batch=B, gemm_mode=mode)
    eng.load_segments(p0)
    x, eps = O.make_batch(2, B, dims)
    loss = float(eng.train_step(dev(x), dev(eps)).item())
    tr = O.Trainer(p0, dims)
    lo, g = tr.step([(x, eps)])
    tol = TOL[mode]
    assert abs(loss - lo[0]) <= tol["loss"] * abs(lo[0])
    got = eng.segments_to_numpy(eng.grads)
    for k in O.SEGMENTS:
        assert rel_l2(got[k], g[k]) < tol["grad"], k
    eng.close()


def test_many_trials_on_one_gpu_equal_sequential_trials(E):
    """T independent trials stepped concurrently (own engine/stream/graph each) end bit-identical to the
    same trials stepped one after another.  train_step_many sizes every trial's kernels for 1/T of the SMs, so the
    sequential twins get the same share (set_sm_share) -- same K-split counts, same summation order."""
    B, T = 128, 3
    xs = [dev(O.make_batch(50 + t, B)[0]) for t in range(T)]
    seq, con = [], []
    for t in range(T):
        for group in (seq, con):
            e = E.VAEEngine(max_batch=B)
            e.load_segments(O.init_params(300 + t))
            e.set_rng(300 + t, 0)
            group.append(e)
        seq[-1].set_sm_share(T)
    for step in range(4):
        for t in range(T):
            seq[t].train_step(xs[(t + step) % T], None, want_loss=False)
        E.train_step_many(con, [xs[(t + step) % T] for t in range(T)])
    for t in range(T):
        seq[t].synchronize(); con[t].synchronize()
        assert torch.equal(seq[t].params, con[t].params)
        assert con[t].state() == (4, 16)
    for e in seq + con:
        e.close()


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("B", [129, 300, 1000])
def test_batches_beyond_one_row_tile_vs_oracle(E, mode, B):
    """Batches that span several 128-row tiles (ragged last tile included): two steps against the oracle, with the
    eps the engine itself drew from Philox fed back to the oracle."""
    tol = TOL[mode]
    p0 = O.init_params(PARAM_SEED)
    eng = make_engine(E, mode, B)
    eng.set_rng(4242, 0)
    tr = O.Trainer(p0)
    keep = {}
    flipped = False
    for t in range(2):
        x, _ = O.make_batch(batch_seed(t) + 7, B)
        loss = float(eng.train_step(dev(x), None).item())
        eps = eng.workspace("eps", B).cpu().numpy()
        L, g, act = O.loss_and_grads(tr.p, x, eps)
        # ReLU is discontinuous: a pre-activation within rounding of 0 can land on different sides on the GPU and in
        # numpy (measured: about one element per 1e5), which changes that hidden unit's gradient row by one sample's
        # worth.  Rows of units whose mask disagrees are left out; a flipped fc3 unit also perturbs everything upstream.
        flip1 = np.flatnonzero(((eng.workspace("h1", B).cpu().numpy() > 0) != (act["h1"] > 0)).any(0))
        flip3 = np.flatnonzero(((eng.workspace("h3", B).cpu().numpy() > 0) != (act["h3"] > 0)).any(0))
        coarse = mode in ("tf32", "bf16")                               # 10- / 8-bit operand mantissas move pre-activations by ~1e-3
        max_flips = 3 if not coarse else p0["W1"].shape[0] * (1 if mode == "tf32" else 3) // 4    # of 400 units x 1000 samples
        assert len(flip1) <= max_flips and len(flip3) <= max_flips, (flip1, flip3)
        flipped = flipped or len(flip1) > 0 or len(flip3) > 0
        for names, fl, n in ((("W1", "b1"), flip1, p0["W1"].shape[0]), (("W3", "b3"), flip3, p0["W3"].shape[0])):
            for k in names:
                keep[k] = np.setdiff1d(keep.get(k, np.arange(n)), fl)
        if t == 0:
            got = eng.segments_to_numpy(eng.grads)
            loose = (50.0 if len(flip3) else 1.0) if not coarse else 3.0
            for k in O.SEGMENTS:
                a, b = (got[k][keep[k]], g[k][keep[k]]) if k in keep else (got[k], g[k])
                lim = tol["grad"] * (loose if k in ("W1", "b1", "W2x", "b2x", "W3", "b3") else 1.0)
                assert rel_l2(a, b) < lim, (k, rel_l2(a, b), flip1, flip3)
        lo, _ = tr.step([(x, eps)])
        assert abs(loss - lo[0]) <= tol["loss"] * abs(lo[0]), (t, loss, lo[0])
    # Adam turns a flipped unit's gradient row into an O(lr) parameter difference: those rows stay excluded
    pN = eng.segments_to_numpy()
    for k in O.SEGMENTS:
        a, b = (pN[k][keep[k]], tr.p[k][keep[k]]) if k in keep else (pN[k], tr.p[k])
        lim = tol["param"] * ((50.0 if flipped else 1.0) if not coarse else 3.0)
        assert rel_l2(a, b) < lim, (k, rel_l2(a, b))
    eng.close()


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("B", [1, 48, 130])
def test_no_stray_writes_into_layout_padding(E, mode, B):
    """compute-sanitizer is closed on this pool, so stray stores are hunted with canaries: the alignment gaps between
    the flat-buffer segments must still be exactly zero after training steps with ragged batches (a tile epilogue that
    ignored its row/column bounds would dirty them), and Adam must keep them zero in params and moments."""
    eng = make_engine(E, mode, B)
    lay, total = O.layout(O.Dims())
    gap = np.ones(total, dtype=bool)
    for off, n in lay.values():
        gap[off:off + n] = False
    assert gap.sum() > 0
    for t in range(3):
        x, _ = O.make_batch(batch_seed(t), B)
        eng.train_step(dev(x), None)
    eng.synchronize()
    for name, buf in (("params", eng.params), ("grads", eng.grads), ("exp_avg", eng.exp_avg), ("exp_avg_sq", eng.exp_avg_sq)):
        assert not buf.cpu().numpy()[gap].any(), name
    eng.close()


@pytest.mark.parametrize("mode", MODES)
def test_submit_wait_pipelined_host_steps_equal_synchronous_ones(E, mode):
    """mdt_train_step_host_submit / _wait with several steps in flight gives, step for step, the losses and the final
    parameters of the synchronous mdt_train_step_host loop (tickets waited out of order too)."""
    B, n = 48, 9
    xs = [torch.from_numpy(O.make_batch(batch_seed(t), B)[0]).pin_memory() for t in range(n)]
    a = make_engine(E, mode, B)
    a.set_rng(99, 0)
    want = [a.train_step_host(x) for x in xs]
    b = make_engine(E, mode, B)
    b.set_rng(99, 0)
    tickets = [b.submit_host(x) for x in xs[:4]]                 # four in flight
    got = {}
    got[1] = b.wait_host(tickets[1])                             # out of order
    got[0] = b.wait_host(tickets[0])
    for t in range(4, n):
        tickets.append(b.submit_host(xs[t]))
        got[t - 2] = b.wait_host(tickets[t - 2])
    for t in range(n - 2, n):
        got[t] = b.wait_host(tickets[t])
    assert [got[t] for t in range(n)] == want
    assert torch.equal(a.params, b.params)
    a.close(); b.close()
