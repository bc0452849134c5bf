"""mdt_loss_bce_kl_noalloc: the loss entry for callers that cannot allocate device memory
(fully committed memory budget, loops that cannot take cudaMalloc's device-wide stall).
Collected last on purpose (file name), so a `-x` run has already exercised everything else
if this fails.  The result must equal the allocating entry point's, call after call, even
though every call reuses one fixed scratch array inside the library."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(seed, B, D=784, Z=20):
    rs = np.random.RandomState(seed)
    f = lambda *s: torch.from_numpy(rs.random_sample(s).astype(np.float32)).cuda()
    n = lambda *s: torch.from_numpy(rs.standard_normal(s).astype(np.float32)).cuda()
    return f(B, D), f(B, D), n(B, Z), n(B, Z)


def test_noalloc_loss_equals_allocating_entry_on_two_calls_in_a_row():
    from multidisttorch_b200 import engine as E
    # two different problems back to back: the second must not see anything the first left in the scratch
    for seed, B in ((1, 128), (2, 37)):
        recon, x, mu, lv = _case(seed, B)
        want_loss, want_dr, want_dm, want_dl = E.loss_bce_kl(recon, x, mu, lv)
        loss = torch.full((1,), float("nan"), device="cuda")
        dr, dm, dl = torch.empty_like(recon), torch.empty_like(mu), torch.empty_like(lv)
        if seed == 1:           # first use on this device resolves the scratch symbol; not part of the allocation check
            E.loss_bce_kl_noalloc(recon, x, mu, lv, loss, dr, dm, dl)
        torch.cuda.synchronize()
        # cudaMemGetInfo sees every device allocation of the process (torch's caching allocator would hide a
        # cudaMalloc made inside the library), torch.cuda.memory_allocated sees torch's own
        free_before, torch_before = torch.cuda.mem_get_info()[0], torch.cuda.memory_allocated()
        E.loss_bce_kl_noalloc(recon, x, mu, lv, loss, dr, dm, dl)
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free_before and torch.cuda.memory_allocated() == torch_before
        w, g = float(want_loss.item()), float(loss.item())
        assert abs(g - w) <= 2e-6 * abs(w), (seed, g, w)          # same tolerance as test_gpu_parity's loss checks
        np.testing.assert_allclose(dr.cpu().numpy(), want_dr.cpu().numpy(), rtol=3e-6, atol=0)
        np.testing.assert_array_equal(dm.cpu().numpy(), want_dm.cpu().numpy())
        np.testing.assert_allclose(dl.cpu().numpy(), want_dl.cpu().numpy(), rtol=3e-6, atol=1.2e-7)


def test_noalloc_rejects_sizes_beyond_the_fixed_scratch():
    from multidisttorch_b200 import engine as E
    B = 1338                                   # 1338*784 = 1,048,992 > 1,048,576
    recon, x, mu, lv = _case(3, B)
    loss = torch.zeros(1, device="cuda")
    with pytest.raises(RuntimeError, match="exceeds the fixed scratch capacity"):
        E.loss_bce_kl_noalloc(recon, x, mu, lv, loss)
    # the largest accepted batch at D=784 still works and agrees
    B = 1337
    recon, x, mu, lv = _case(4, B)
    E.loss_bce_kl_noalloc(recon, x, mu, lv, loss)
    want = float(E.loss_bce_kl(recon, x, mu, lv, want_grads=False)[0].item())
    assert abs(float(loss.item()) - want) <= 2e-6 * abs(want)


def test_noalloc_loss_against_the_oracle_and_the_reference_corners():
    """Not only equal to its sibling entry point: checked against the CPU oracle on random data and against the
    reference-generated clamp / saturation corners (tests/golden/loss_edge.npz)."""
    from multidisttorch_b200 import engine as E
    from oracle import vae_numpy as O
    from tests.golden_cases import load, rel_l2
    recon, x, mu, lv = _case(9, 200)
    loss = torch.zeros(1, device="cuda")
    dr, dm, dl = torch.empty_like(recon), torch.empty_like(mu), torch.empty_like(lv)
    E.loss_bce_kl_noalloc(recon, x, mu, lv, loss, dr, dm, dl)
    r, xx, m, l = (t.cpu().numpy() for t in (recon, x, mu, lv))
    ref = O.loss_function(r, xx, m, l)
    rdr, rdm, rdl = O.loss_grads(r, xx, m, l)
    assert abs(float(loss.item()) - ref) <= 2e-6 * abs(ref)
    assert rel_l2(dr.cpu().numpy(), rdr) < 1e-6 and rel_l2(dl.cpu().numpy(), rdl) < 1e-6
    np.testing.assert_array_equal(dm.cpu().numpy(), rdm)
    g = load("loss_edge")
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    recon, x, mu, lv = d(g["recon"]), d(g["x"]), d(g["mu"]), d(g["logvar"])
    dr, dm, dl = torch.empty_like(recon), torch.empty_like(mu), torch.empty_like(lv)
    E.loss_bce_kl_noalloc(recon, x, mu, lv, loss, dr, dm, dl)
    assert abs(float(loss.item()) - float(g["loss"])) <= 2e-6 * abs(float(g["loss"]))
    np.testing.assert_allclose(dr.cpu().numpy(), g["d_recon"], rtol=3e-6, atol=0)
    np.testing.assert_array_equal(dm.cpu().numpy(), g["d_mu"])
    np.testing.assert_allclose(dl.cpu().numpy(), g["d_logvar"], rtol=3e-6, atol=1.2e-7)
