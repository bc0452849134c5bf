"""Compat mode on the GPU: the torch names the reference scripts call
(DistributedDataParallel / optim.Adam, ref:vae-hpo.py:130-131) are replaced by the
engine-backed objects, and the ref:vae-hpo.py:69-74 loop written against them matches
plain torch CUDA running the same model with the same generator state."""
import pytest
import torch
import torch.nn.functional as F

from oracle import torch_port as TP
from oracle import vae_numpy as O
from tests.golden_cases import rel_l2

pytestmark = pytest.mark.gpu


def ref_loss(recon, x, mu, logvar):        # same calls as ref:vae-hpo.py:49-58
    bce = F.binary_cross_entropy(recon, x.view(-1, 784), reduction="sum")
    return bce - 0.5 * torch.sum(1 + logvar - mu.pow(2) - logvar.exp())


def loop(model, opt, xs):
    losses = []
    model.train()
    for x in xs:
        opt.zero_grad()
        recon, mu, logvar = model(x)
        loss = ref_loss(recon, x, mu, logvar)
        loss.backward()
        losses.append(loss.item())
        opt.step()
    return losses


def test_dropin_training_loop_matches_torch_cuda():
    from multidisttorch_b200 import dropin
    torch.backends.cuda.matmul.allow_tf32 = False
    xs = [torch.rand(128, 1, 28, 28, generator=torch.Generator().manual_seed(i)).cuda() for i in range(6)]
    p0 = O.init_params(21)

    plain = TP.VAEPort().cuda()
    TP.load_segments(plain, p0)
    torch.manual_seed(77)
    want = loop(plain, torch.optim.Adam(plain.parameters(), lr=1e-3), xs)

    assert dropin.install() or dropin._ORIG
    try:
        m = TP.VAEPort().cuda()
        TP.load_segments(m, p0)
        torch.manual_seed(77)
        wrapped = torch.nn.parallel.DistributedDataParallel(m, process_group=None)
        assert isinstance(wrapped, dropin.EngineModule)
        opt = torch.optim.Adam(wrapped.parameters(), lr=1e-3)
        assert isinstance(opt, dropin.EngineAdam)
        got = loop(wrapped, opt, xs)
        # eval / no_grad path and decode through the original module (ref:vae-hpo.py:101-105,165)
        wrapped.eval()
        with torch.no_grad():
            recon, mu, logvar = wrapped(xs[0])
            z = torch.randn(64, 20).cuda()
            dec = wrapped.module.decode(z)
        plain.eval()
        assert rel_l2(dec.cpu().numpy(), plain.decode(z).detach().cpu().numpy()) < 1e-4
        # anything that is not the reference VAE falls through to the real torch classes
        lin = torch.nn.Linear(4, 4).cuda()
        o2 = torch.optim.Adam(lin.parameters(), lr=1e-2)
        assert isinstance(o2, dropin._ORIG["adam"]) and not isinstance(o2, dropin.EngineAdam)
    finally:
        dropin.uninstall()
    for a, b in zip(got, want):
        assert abs(a - b) <= 2e-5 * abs(b), (got, want)
    for (n, p), (_, q) in zip(wrapped.module.named_parameters(), plain.named_parameters()):
        assert rel_l2(p.detach().cpu().numpy(), q.detach().cpu().numpy()) < 2e-4, n


def test_dropin_adam_takes_per_trial_hyperparameters_from_env(monkeypatch):
    """with MDT_HPO_* set, the optimizer that stands in for `optim.Adam(model.parameters(), lr=1e-3)`
    trains with the trial's own lr / betas, and the result equals plain torch trained with those values."""
    from multidisttorch_b200 import dropin
    torch.backends.cuda.matmul.allow_tf32 = False
    monkeypatch.setenv("MDT_HPO_LR", "3e-4")
    monkeypatch.setenv("MDT_HPO_BETA1", "0.8")
    xs = [torch.rand(64, 1, 28, 28, generator=torch.Generator().manual_seed(100 + i)).cuda() for i in range(5)]
    p0 = O.init_params(22)
    plain = TP.VAEPort().cuda()
    TP.load_segments(plain, p0)
    torch.manual_seed(78)
    want = loop(plain, torch.optim.Adam(plain.parameters(), lr=3e-4, betas=(0.8, 0.999)), xs)
    assert dropin.install() or dropin._ORIG
    try:
        m = TP.VAEPort().cuda()
        TP.load_segments(m, p0)
        torch.manual_seed(78)
        wrapped = torch.nn.parallel.DistributedDataParallel(m, process_group=None)
        opt = torch.optim.Adam(wrapped.parameters(), lr=1e-3)          # what the script asks for
        assert isinstance(opt, dropin.EngineAdam)
        assert opt.defaults["lr"] == 3e-4 and opt.defaults["betas"] == (0.8, 0.999)
        got = loop(wrapped, opt, xs)
    finally:
        dropin.uninstall()
    for a, b in zip(got, want):
        assert abs(a - b) <= 2e-5 * abs(b), (got, want)
    for (n, p), (_, q) in zip(wrapped.module.named_parameters(), plain.named_parameters()):
        assert rel_l2(p.detach().cpu().numpy(), q.detach().cpu().numpy()) < 2e-4, n


def test_dropin_leaves_lookalike_models_to_torch(capsys):
    """A model with the reference VAE's layer names and shapes but another forward (here: no sigmoid, LeakyReLU) must
    NOT be swapped for the reference VAE's math: the shim checks behaviour on a small batch and falls through."""
    from multidisttorch_b200 import dropin

    class LookAlike(TP.VAEPort):
        def forward(self, x, eps=None):
            h1 = F.leaky_relu(self.fc1(x.view(-1, self.in_dim)), 0.2)
            mu, logvar = self.fc21(h1), self.fc22(h1)
            z = mu + torch.randn_like(logvar) * torch.exp(0.5 * logvar)
            return self.fc4(F.leaky_relu(self.fc3(z), 0.2)), mu, logvar

    assert dropin.install() or dropin._ORIG
    try:
        m = LookAlike().cuda()
        with pytest.raises(Exception, match="process group|init_process_group"):   # the real DDP constructor ran
            torch.nn.parallel.DistributedDataParallel(m, process_group=None)
        assert "left to torch DDP" in capsys.readouterr().out
        ok = torch.nn.parallel.DistributedDataParallel(TP.VAEPort().cuda(), process_group=None)
        assert isinstance(ok, dropin.EngineModule)
        assert "-> B200 engine" in capsys.readouterr().out
    finally:
        dropin.uninstall()


@pytest.mark.skipif(__import__("os").environ.get("MDT_TEST_UNVALIDATED") != "1",
                    reason="written after the round's GPU budget was spent: never run on a GPU yet (set MDT_TEST_UNVALIDATED=1); "
                           "the SGD kernel itself is covered by tests/test_gpu_steps.py, the shim logic by tests/test_dropin_shims.py")
def test_dropin_sgd_training_loop_matches_torch_cuda():
    """north_star (c): `optim.SGD(model.parameters(), ...)` on the wrapped model steps through the engine's flat SGD kernel
    and follows plain torch-CUDA SGD (momentum, weight decay, Nesterov) on the same model and generator state."""
    from multidisttorch_b200 import dropin
    torch.backends.cuda.matmul.allow_tf32 = False
    xs = [torch.rand(128, 1, 28, 28, generator=torch.Generator().manual_seed(40 + i)).cuda() for i in range(6)]
    p0 = O.init_params(23)
    kw = dict(lr=2e-6, momentum=0.9, weight_decay=1e-2, nesterov=True)    # the loss is a SUM over 128 x 784 pixels

    plain = TP.VAEPort().cuda()
    TP.load_segments(plain, p0)
    start = {n: p.detach().clone() for n, p in plain.named_parameters()}
    torch.manual_seed(79)
    want = loop(plain, torch.optim.SGD(plain.parameters(), **kw), xs)

    assert dropin.install() or dropin._ORIG
    try:
        m = TP.VAEPort().cuda()
        TP.load_segments(m, p0)
        torch.manual_seed(79)
        wrapped = torch.nn.parallel.DistributedDataParallel(m, process_group=None)
        opt = torch.optim.SGD(wrapped.parameters(), **kw)
        assert isinstance(opt, dropin.EngineSGD)
        got = loop(wrapped, opt, xs)
        lin = torch.nn.Linear(4, 4).cuda()
        assert not isinstance(torch.optim.SGD(lin.parameters(), lr=1e-2), dropin.EngineSGD)
    finally:
        dropin.uninstall()
    for a, b in zip(got, want):
        assert abs(a - b) <= 2e-5 * abs(b), (got, want)
    for (n, p), (_, q) in zip(wrapped.module.named_parameters(), plain.named_parameters()):
        moved = (q.detach() - start[n]).cpu().numpy()
        assert rel_l2((p.detach() - start[n]).cpu().numpy(), moved) < 1e-3, n      # the UPDATE, not the parameter
