"""Host logic of the optimizer shims (multidisttorch_b200/dropin.py) without a GPU: which constructor calls are handed
to the engine-backed optimizer and which fall through to the real torch class.  The engine itself is a recording stub
here -- the arithmetic is checked on the GPU (tests/test_gpu_steps.py for the SGD kernel, tests/test_gpu_dropin.py)."""
import pytest
import torch
from torch import nn

from multidisttorch_b200 import dropin


class _StubEngine:
    def __init__(self):
        self.calls = []
        self.exp_avg = torch.ones(8)

    def set_sgd(self, *a, **k):
        self.calls.append(("set_sgd", a, k))

    def set_adam(self, *a, **k):
        self.calls.append(("set_adam", a, k))

    def adam_step(self):
        self.calls.append(("step",))

    def state(self):
        return 3, 0


class _StubOwner:
    """What EngineSGD / EngineAdam need from an EngineModule."""

    def __init__(self):
        self.module = nn.ModuleList([nn.Linear(2, 2) for _ in range(5)])     # 10 parameters, like the VAE
        self.engine = _StubEngine()
        self.process_group = None
        for p in self.module.parameters():
            dropin._ENGINE_OF_PARAM[id(p)] = self

    def forget(self):
        for p in self.module.parameters():
            dropin._ENGINE_OF_PARAM.pop(id(p), None)


@pytest.fixture
def owner():
    o = _StubOwner()
    yield o
    o.forget()


def test_sgd_shim_falls_through_for_ordinary_parameters():
    SGD = dropin._make_sgd_shim(torch.optim.SGD)
    lin = nn.Linear(3, 1)
    opt = SGD(lin.parameters(), lr=0.5, momentum=0.9)            # a generator argument, as scripts pass it
    assert type(opt) is SGD and isinstance(opt, torch.optim.SGD) and not isinstance(opt, dropin.EngineSGD)
    w0 = lin.weight.detach().clone()
    lin(torch.ones(4, 3)).sum().backward()
    opt.step()
    assert torch.allclose(lin.weight, w0 - 0.5 * 4.0)             # d/dw sum(w.x + b) over four rows of ones = 4
    assert opt.param_groups[0]["momentum"] == 0.9
    with pytest.raises(ValueError):                               # torch's own argument checks still run
        SGD(lin.parameters(), lr=0.1, nesterov=True)


def test_sgd_shim_hands_the_engines_parameters_to_the_engine(owner):
    SGD = dropin._make_sgd_shim(torch.optim.SGD)
    opt = SGD(owner.module.parameters(), lr=1e-4, momentum=0.9, weight_decay=1e-3, nesterov=True)
    assert isinstance(opt, dropin.EngineSGD)
    assert owner.engine.calls == [("set_sgd", (1e-4, 0.9, 0.0, 1e-3, True), {"completed_steps": 0})]
    assert float(owner.engine.exp_avg.abs().sum()) == 0.0         # a new optimizer starts without a momentum buffer
    assert owner._sgd_h == (1e-4, 0.9, 0.0, 1e-3, True)
    opt.zero_grad()
    opt.step()
    assert owner.engine.calls[-1] == ("step",)
    sd = opt.state_dict()
    assert sd["step"] == 3 and sd["param_groups"][0]["lr"] == 1e-4 and "params" not in sd["param_groups"][0]
    assert len(opt.param_groups[0]["params"]) == 10


def test_sgd_shim_leaves_unusual_requests_to_torch(owner):
    SGD = dropin._make_sgd_shim(torch.optim.SGD)
    plist = list(owner.module.parameters())
    assert not isinstance(SGD(plist[:9], lr=1e-3), dropin.EngineSGD)                      # not the whole model
    assert not isinstance(SGD(plist, lr=1e-3, maximize=True), dropin.EngineSGD)           # an option the kernel lacks
    assert not isinstance(SGD([{"params": plist}], lr=1e-3), dropin.EngineSGD)            # parameter groups
    other = nn.Linear(2, 2)
    assert not isinstance(SGD(plist[:8] + list(other.parameters()), lr=1e-3), dropin.EngineSGD)   # two owners
    assert owner.engine.calls == []


def test_adam_after_sgd_clears_the_sgd_choice(owner):
    opt = dropin.EngineSGD(owner, 1e-3, 0.0, 0.0, 0.0, False)
    assert owner._sgd_h is not None
    dropin.EngineAdam(owner, 1e-3, (0.9, 0.999), 1e-8)
    assert owner._sgd_h is None and owner._adam_h == (1e-3, (0.9, 0.999), 1e-8)
    assert owner.engine.calls[-1][0] == "set_adam"
    del opt
