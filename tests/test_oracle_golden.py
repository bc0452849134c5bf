"""The numpy oracle against the golden vectors the real reference produced
(oracle/make_golden.py).  CPU only."""
import numpy as np
import pytest

from oracle import vae_numpy as O
from tests.golden_cases import PARAM_SEED, batch_seed, check_fp, load, rel_l2

# oracle (numpy fp32) vs reference (torch fp32 CPU): differences are BLAS summation order only
TOL_ACT = 1e-5
TOL_GRAD = 2e-5
TOL_PARAM = 1e-5
TOL_LOSS = 2e-6


@pytest.mark.parametrize("tag", ["step_b128", "step_b48", "step_b1", "step_b512", "step_h4096_b128"])
def test_train_steps_match_reference(tag):
    g = load(tag)
    B, n = int(g["B"]), int(g["n_steps"])
    dims = O.Dims(H=int(g["H"]) if "H" in g else 400)
    # K = 4096 dot products: numpy's and torch's summation orders differ by ~1e-6, which Adam's first steps
    # amplify to 6.9e-5 on W1 (measured by oracle/make_golden.py)
    TOL_PARAM = 1e-5 if dims.H == 400 else 2e-4
    tr = O.Trainer(O.init_params(PARAM_SEED, dims), dims)
    for t in range(n):
        x, eps = O.make_batch(batch_seed(t), B, dims)
        if t == 0:
            L, grads, act = O.loss_and_grads(tr.p, x, eps, dims)
            for k_ref, k in (("recon", "recon"), ("mu", "mu"), ("logvar", "lv")):
                check_fp(g, "step0", k_ref, act[k], TOL_ACT, tag)
            for k in O.SEGMENTS:
                check_fp(g, "grad0", k, grads[k], TOL_GRAD, tag)
        losses, _ = tr.step([(x, eps)])
        assert abs(losses[0] - g["losses"][t]) <= TOL_LOSS * abs(g["losses"][t]), (tag, t)
    for k in O.SEGMENTS:
        check_fp(g, "paramN", k, tr.p[k], TOL_PARAM, tag)
        check_fp(g, "mN", k, tr.m[k], max(TOL_GRAD, TOL_PARAM), tag)
        check_fp(g, "vN", k, tr.v[k], 2 * max(TOL_GRAD, TOL_PARAM), tag)


def test_eval_forward_and_decode():
    g = load("eval_b64")
    p = O.init_params(PARAM_SEED)
    x, eps = O.make_batch(batch_seed(77), int(g["B"]))
    act = O.forward(p, x, eps)
    for k_ref, k in (("recon", "recon"), ("mu", "mu"), ("logvar", "lv")):
        check_fp(g, "eval", k_ref, act[k], TOL_ACT)
    L = O.loss_function(act["recon"], x, act["mu"], act["lv"])
    assert abs(L - float(g["loss"])) <= TOL_LOSS * abs(float(g["loss"]))
    zs = np.random.RandomState(99).standard_normal((64, 20)).astype(np.float32)
    h3 = np.maximum(zs @ p["W3"].T + p["b3"], 0)
    dec = O.sigmoid(h3 @ p["W4"].T + p["b4"])
    check_fp(g, "eval", "decode64", dec, TOL_ACT)


def test_loss_clamp_and_saturation_corners():
    g = load("loss_edge")
    b, k = O.loss_terms(g["recon"], g["x"], g["mu"], g["logvar"])
    assert abs((b + k) - float(g["loss"])) <= 1e-6 * abs(float(g["loss"]))
    dr, dm, dl = O.loss_grads(g["recon"], g["x"], g["mu"], g["logvar"])
    assert np.all(np.isfinite(dr))
    # first three rows hold recon = 0, 1, denormals ... : every element must agree, not just the norm
    np.testing.assert_allclose(dr[:3], g["d_recon"][:3], rtol=2e-6, atol=0)
    assert rel_l2(dr, g["d_recon"]) < 1e-6
    np.testing.assert_array_equal(dm, g["d_mu"])
    np.testing.assert_allclose(dl, g["d_logvar"], rtol=2e-6, atol=1.2e-7)  # |0.5(e^lv-1)| cancels near lv=0


def test_step_with_logits_on_the_bce_corners():
    from oracle.make_golden import EDGE_LOGITS, edge_setup
    g = load("step_edge_b8")
    B, n = int(g["B"]), len(EDGE_LOGITS)
    x, eps = O.make_batch(batch_seed(0), B)
    p0, x = edge_setup(O.init_params(PARAM_SEED), x)
    L, grads, act = O.loss_and_grads(p0, x, eps)
    assert abs(L - g["losses"][0]) <= TOL_LOSS * abs(g["losses"][0])
    np.testing.assert_allclose(act["recon"][:3, :n], g["recon_edge"], rtol=1e-6, atol=1e-44)
    np.testing.assert_allclose(grads["b4"][:n], g["grad0.b4.edge"], rtol=2e-5, atol=1e-30)
    for k in O.SEGMENTS:
        check_fp(g, "grad0", k, grads[k], TOL_GRAD, "edge")


@pytest.mark.parametrize("tag", ["ddp_1x2_b32", "ddp_2x2_b16", "ddp_1x4_b32", "ddp_1x8_b16", "ddp_2x4_b128"])
def test_ddp_groups_match_reference(tag):
    g = load(tag)
    world, ng, B, n = (int(g[k]) for k in ("world", "ngroups", "B", "n_steps"))
    shift = int(g["seed_shift"]) if "seed_shift" in g else 0
    k = world // ng
    for gid in range(ng):
        members = list(range(k * gid, k * gid + k))
        tr = O.Trainer(O.init_params(PARAM_SEED + gid))
        for t in range(n):
            losses, _ = tr.step([O.make_batch(batch_seed(t, r) + shift, B) for r in members])
            for i, r in enumerate(members):
                ref = float(g[f"losses.rank{r}"][t])
                assert abs(losses[i] - ref) <= TOL_LOSS * abs(ref), (tag, gid, t, r)
        for name in O.SEGMENTS:
            check_fp(g, f"g{gid}.paramN", name, tr.p[name], TOL_PARAM, tag)


def test_pack_unpack_layout():
    d = O.Dims()
    lay, total = O.layout(d)
    assert d.n_params() == 652824
    assert all(off % O.ALIGN == 0 for off, _ in lay.values())
    p = O.init_params(3)
    flat = O.pack(p, d)
    assert flat.size == total
    q = O.unpack(flat, d)
    for k in O.SEGMENTS:
        np.testing.assert_array_equal(p[k], q[k])
    d2 = O.Dims(H=4096)
    assert d2.n_params() == 6677304


def test_philox_restatement_known_answer():
    """Philox4x32-10 known-answer vectors from the Random123 distribution
    (kat_vectors): counter/key all-zero and all-ones."""
    z = O.philox4x32_10(np.zeros((1, 4), np.uint32), np.zeros(2, np.uint32))[0]
    assert [hex(int(v)) for v in z] == ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]
    f = np.uint32(0xFFFFFFFF)
    o = O.philox4x32_10(np.full((1, 4), f, np.uint32), np.full(2, f, np.uint32))[0]
    assert [hex(int(v)) for v in o] == ["0x408f276d", "0x41c83b0e", "0xa20bc7c6", "0x6d5451fd"]
    e = O.cuda_randn(1234, 0, 2560)
    assert e.dtype == np.float32 and abs(float(e.mean())) < 0.08 and 0.9 < float(e.std()) < 1.1
