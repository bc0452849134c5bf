"""mdt_train_steps (the inner loop of train(), ref:vae-hpo.py:67-74, as graphs of several steps over a device-resident
batch source), the uint8 dataset source (gather + /255 inside the step, ref:vae-hpo.py:135-150) and the SGD update."""
import numpy as np
import pytest
import torch

from oracle import vae_numpy as O
from tests.golden_cases import PARAM_SEED, batch_seed, rel_l2

pytestmark = pytest.mark.gpu


def pair(E, mode, B, **kw):
    engs = []
    for _ in range(2):
        e = E.VAEEngine(max_batch=B, gemm_mode=mode, **kw)
        e.load_segments(O.init_params(PARAM_SEED))
        e.set_rng(31, 0)
        engs.append(e)
    return engs


@pytest.fixture(scope="module")
def E():
    from multidisttorch_b200 import engine
    return engine


@pytest.mark.parametrize("mode", ["fp32", "tf32x3", "tf32", "bf16"])
@pytest.mark.parametrize("B", [128, 48])
def test_train_steps_over_a_pool_equal_single_steps_bit_for_bit(E, mode, B):
    a, b = pair(E, mode, B)
    pool = torch.rand(7, B, 784, generator=torch.Generator().manual_seed(B)).cuda()
    n = 27                                              # three 8-step graphs + 3 single steps, wrapping the pool
    losses = torch.zeros(n, device="cuda")
    a.train_steps(pool, 5, n, losses)
    want = [float(b.train_step(pool[(5 + i) % 7], None).item()) for i in range(n)]
    a.synchronize()
    assert losses.cpu().tolist() == want
    assert torch.equal(a.params, b.params) and torch.equal(a.exp_avg_sq, b.exp_avg_sq)
    assert a.state() == b.state()
    # a second call that continues where the first stopped, and one that jumps
    a.train_steps(pool, (5 + n) % 7, 9, None)
    a.train_steps(pool, 2, 2, None)
    for i in list(range(n, n + 9)):
        b.train_step(pool[(5 + i) % 7], None)
    for i in (2, 3):
        b.train_step(pool[i], None)
    a.synchronize(); b.synchronize()
    assert torch.equal(a.params, b.params)
    a.close(); b.close()


@pytest.mark.parametrize("mode", ["tf32x3", "bf16"])
def test_train_steps_from_uint8_images_equal_the_torch_pipeline(E, mode):
    """The step's first kernel gathers rows of a uint8 image table through the sampler's index order and divides by
    255 (what datasets.MNIST + ToTensor + DistributedSampler + DataLoader hand to ref:vae-hpo.py:67-68)."""
    B, n_img = 64, 1000
    rs = np.random.RandomState(5)
    images = torch.from_numpy(rs.randint(0, 256, (n_img, 784), dtype=np.uint8)).cuda()
    order = torch.from_numpy(rs.permutation(n_img)[: 10 * B].astype(np.int64)).cuda()
    a, b = pair(E, mode, B)
    losses = torch.zeros(10, device="cuda")
    a.train_steps_u8(images, order, B, 0, 10, losses)
    want = []
    for i in range(10):
        # ToTensor(): uint8 -> float32, then a true division by 255 ON THE CPU (torch's CUDA kernel for a division by a
        # scalar multiplies by the reciprocal instead, which differs in the last bit)
        x = (images[order[i * B:(i + 1) * B]].cpu().to(torch.float32) / 255).cuda()
        want.append(float(b.train_step(x.contiguous(), None).item()))
    a.synchronize()
    assert losses.cpu().tolist() == want and torch.equal(a.params, b.params)
    a.close(); b.close()


@pytest.mark.parametrize("mode", ["fp32", "tf32x3"])
# (the loss is a SUM over the batch, so its gradients are O(100): learning rates that keep SGD stable are ~1e-6)
@pytest.mark.parametrize("kw", [dict(lr=2e-6), dict(lr=1e-6, momentum=0.9), dict(lr=1e-6, momentum=0.9, nesterov=True, weight_decay=1e-3),
                                dict(lr=1e-6, momentum=0.8, dampening=0.2)], ids=lambda k: "-".join(f"{a}{b}" for a, b in k.items()))
def test_sgd_steps_match_the_oracle(E, mode, kw):
    """north_star (c): SGD in place of Adam.  The oracle's SGD is pinned against torch.optim.SGD in
    tests/test_hpo_surface.py; here the engine against the oracle over 4 steps (momentum buffer included)."""
    B = 48
    p0 = O.init_params(PARAM_SEED)
    eng = E.VAEEngine(max_batch=B, gemm_mode=mode)
    eng.load_segments(p0)
    eng.set_sgd(**kw)
    full = dict(dict(momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False), **kw)
    tr = O.Trainer(p0, sgd=full)
    for t in range(4):
        x, eps = O.make_batch(batch_seed(t), B)
        loss = float(eng.train_step(torch.from_numpy(x).cuda(), torch.from_numpy(eps).cuda()).item())
        lo, _ = tr.step([(x, eps)])
        assert abs(loss - lo[0]) <= 1e-5 * abs(lo[0]), (t, loss, lo[0])
    p = eng.segments_to_numpy()
    m = eng.segments_to_numpy(eng.exp_avg)
    for k in O.SEGMENTS:
        assert rel_l2(p[k], tr.p[k]) < 2e-5, (k, rel_l2(p[k], tr.p[k]))
        if full["momentum"]:
            assert rel_l2(m[k], tr.m[k]) < 2e-4, (k, rel_l2(m[k], tr.m[k]))
    assert eng.state()[0] == 4
    eng.close()


@pytest.mark.parametrize("mode", ["tf32x3", "bf16"])
def test_host_chunks_equal_synchronous_host_steps(E, mode):
    """mdt_train_steps_host_submit: chunks of host batches (ragged last chunk, two chunks in flight) give, step for step,
    the losses and final parameters of the synchronous mdt_train_step_host loop."""
    B, n = 48, 41
    xs = torch.rand(n, B, 784, generator=torch.Generator().manual_seed(8)).pin_memory()
    a, b = pair(E, mode, B)
    want = [a.train_step_host(xs[i]) for i in range(n)]
    tickets = []
    got = []
    for lo in range(0, n, 16):
        tickets.append(b.submit_host_chunk(xs[lo:lo + 16]))
        if len(tickets) == 2:                           # keep one chunk in flight ahead
            got += [b.wait_host(t) for t in tickets.pop(0)]
    for ch in tickets:
        got += [b.wait_host(t) for t in ch]
    assert got == want
    assert torch.equal(a.params, b.params) and a.state() == b.state()
    a.close(); b.close()


@pytest.mark.parametrize("mode", ["tf32x3", "bf16"])
@pytest.mark.parametrize("T", [2, 3])
def test_grouped_trials_equal_trials_stepped_alone(E, mode, T):
    """T trials stepped with ONE launch per layer (mdt_group_train_steps) end bit-identical to the same
    trials stepped alone with their kernels sized for 1/T of the SMs, and keep their own losses / Philox streams."""
    B = 128
    pools = [torch.rand(5, B, 784, generator=torch.Generator().manual_seed(40 + t)).cuda() for t in range(T)]
    grouped, alone = [], []
    for t in range(T):
        for lst in (grouped, alone):
            e = E.VAEEngine(max_batch=B, gemm_mode=mode)
            e.load_segments(O.init_params(500 + t))
            e.set_rng(600 + t, 0)
            lst.append(e)
        alone[-1].set_sm_share(T)
    grp = E.TrialGroup(grouped)
    n = 21                                              # one 16-step graph + 5 single steps
    grp.train_steps(pools, 2, n)
    grp.train_steps(pools, (2 + n) % 5, 3)              # continues where it stopped
    grp.synchronize()
    for t in range(T):
        alone[t].train_steps(pools[t], 2, n + 3)
        alone[t].synchronize(); grouped[t].synchronize()
        assert torch.equal(grouped[t].params, alone[t].params), t
        assert torch.equal(grouped[t].exp_avg_sq, alone[t].exp_avg_sq), t
        assert grouped[t].state() == alone[t].state() == (n + 3, 4 * (n + 3))
    # the engines of a group remain usable on their own afterwards
    l0 = float(grouped[0].train_step(pools[0][0], None).item())
    l1 = float(alone[0].train_step(pools[0][0], None).item())
    alone[0].set_sm_share(1)
    assert np.isfinite(l0) and abs(l0 - l1) <= 1e-5 * abs(l1)
    grp.close()
    for e in grouped + alone:
        e.close()


@pytest.mark.parametrize("mode", ["tf32x3", "bf16"])
def test_train_steps_with_a_separate_optimizer_kernel_equal_single_steps(E, mode, monkeypatch):
    """Steps whose optimizer is its own kernel (what DDP groups and SGD run; forced here on one GPU with
    MDT_NO_FUSED_ADAM=1): inside a multi-step graph the refresh of the transposed weight shadows is moved to a side
    stream and joined before the next step's first reader.  Must stay bit-identical to single steps."""
    monkeypatch.setenv("MDT_NO_FUSED_ADAM", "1")
    B = 64
    a, b = pair(E, mode, B)
    pool = torch.rand(6, B, 784, generator=torch.Generator().manual_seed(77)).cuda()
    n = 37
    losses = torch.zeros(n, device="cuda")
    a.train_steps(pool, 1, n, losses)
    want = [float(b.train_step(pool[(1 + i) % 6], None).item()) for i in range(n)]
    a.synchronize()
    assert losses.cpu().tolist() == want
    assert torch.equal(a.params, b.params) and torch.equal(a.exp_avg, b.exp_avg)
    a.close(); b.close()
