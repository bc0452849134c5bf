"""per-trial hyper-parameter surface (multidisttorch_b200/hpo.py).  CPU part: parsing, trial ids of
the subgroups, and the oracle's Adam against torch.optim.Adam for NON-default hyper-parameters (the golden vectors
only pin lr=1e-3 / default betas).  GPU part: the engine with per-trial hyper-parameters against the oracle."""
import numpy as np
import pytest
import torch

from multidisttorch_b200 import hpo
from oracle import vae_numpy as O
from tests import dist_launch
from tests.golden_cases import PARAM_SEED, batch_seed, rel_l2


def test_trial_config_lists_wrap_and_defaults():
    env = {"MDT_HPO_LR": "1e-3, 3e-4,1e-2", "MDT_HPO_BETA1": "0.8", "MDT_HPO_BATCH": "128,256"}
    c0, c1, c4 = (hpo.trial_config(g, env=env, lr=5e-3) for g in (0, 1, 4))
    assert (c0.lr, c1.lr, c4.lr) == (1e-3, 3e-4, 3e-4)            # 4 % 3 == 1
    assert c0.beta1 == c1.beta1 == 0.8 and c0.beta2 == 0.999 and c0.eps == 1e-8
    assert (c0.batch_size, c1.batch_size, c4.batch_size) == (128, 256, 128)
    assert c1.betas == (0.8, 0.999)
    assert hpo.trial_config(None, env={}, lr=5e-3).lr == 5e-3    # nothing set: the script's own value stands
    assert hpo.overridden(env) and not hpo.overridden({})


@pytest.mark.parametrize("env", [{"MDT_HPO_LR": "abc"}, {"MDT_HPO_LR": "-1"}, {"MDT_HPO_BETA2": "1.0"},
                                 {"MDT_HPO_BATCH": "0"}, {"MDT_HPO_EPS": ","}])
def test_trial_config_rejects_bad_values(env):
    with pytest.raises(ValueError):
        hpo.trial_config(0, env=env)


def test_group_index_of_subgroups_over_gloo():
    world, ngroups = 4, 2
    out, codes, logs = dist_launch.launch(world, ["plain", ngroups, "{out}"])
    assert codes == [0] * world, logs
    res = dist_launch.results(out, world)
    assert all(r["hpo_group_index"] == [0, 1] for r in res)          # every rank knows every group's trial id
    assert [r["hpo_my_group"] for r in res] == [0, 0, 1, 1]


@pytest.mark.parametrize("lr,b1,b2,eps", [(3e-4, 0.8, 0.95, 1e-6), (1e-2, 0.0, 0.9, 1e-8)])
def test_oracle_adam_matches_torch_for_non_default_hyperparameters(lr, b1, b2, eps):
    rng = np.random.RandomState(7)
    p0 = rng.randn(257).astype(np.float32)
    p = p0.copy(); m = np.zeros_like(p); v = np.zeros_like(p)
    tp = torch.nn.Parameter(torch.from_numpy(p0.copy()))
    opt = torch.optim.Adam([tp], lr=lr, betas=(b1, b2), eps=eps)
    for t in range(1, 8):
        g = rng.randn(257).astype(np.float32)
        O.adam_update(p, g, m, v, t, lr=lr, b1=b1, b2=b2, eps=eps)
        tp.grad = torch.from_numpy(g.copy())
        opt.step()
    assert rel_l2(p, tp.detach().numpy()) < 2e-7


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["fp32", "tf32x3"])
def test_engine_with_per_trial_hyperparameters_vs_oracle(mode):
    """Two trials with different lr / betas / eps / batch (what MDT_HPO_* would hand them) on one GPU."""
    from multidisttorch_b200 import engine as E
    env = {"MDT_HPO_LR": "3e-4,1e-2", "MDT_HPO_BETA1": "0.8,0.5", "MDT_HPO_BETA2": "0.95", "MDT_HPO_EPS": "1e-6,1e-8",
           "MDT_HPO_BATCH": "48,128"}
    for gid in (0, 1):
        cfg = hpo.trial_config(gid, env=env)
        B = cfg.batch_size
        eng = E.VAEEngine(max_batch=B, gemm_mode=mode, lr=cfg.lr, betas=cfg.betas, eps=cfg.eps)
        eng.load_segments(O.init_params(PARAM_SEED + gid))
        tr = O.Trainer(O.init_params(PARAM_SEED + gid), lr=cfg.lr, b1=cfg.beta1, b2=cfg.beta2, eps=cfg.eps)
        for t in range(6):
            x, eps = O.make_batch(batch_seed(t, gid), B)
            loss = float(eng.train_step(torch.from_numpy(x).cuda(), torch.from_numpy(eps).cuda()).item())
            ref_loss, _ = tr.step([(x, eps)])
            assert abs(loss - ref_loss[0]) <= (1e-5 if mode == "fp32" else 2e-5) * abs(ref_loss[0]), (gid, t, loss, ref_loss)
        got = eng.segments_to_numpy()
        tol = 1e-4 if mode == "fp32" else 2e-3      # larger lr amplifies the split-TF32 gradient differences (§2)
        for k in O.SEGMENTS:
            assert rel_l2(got[k], tr.p[k]) < tol, (gid, k, rel_l2(got[k], tr.p[k]))
        assert eng.state()[0] == 6
        eng.close()


@pytest.mark.parametrize("kw", [dict(lr=0.05), dict(lr=0.01, momentum=0.9), dict(lr=0.01, momentum=0.9, dampening=0.1),
                                dict(lr=0.02, momentum=0.8, nesterov=True), dict(lr=0.01, momentum=0.9, weight_decay=1e-2)],
                         ids=lambda k: "-".join(f"{a}{b}" for a, b in k.items()))
def test_oracle_sgd_is_torch_sgd(kw):
    """north_star (c) "fused per-trial Adam/SGD update": the oracle's SGD restatement pinned against torch.optim.SGD."""
    import numpy as np
    import torch
    from oracle import vae_numpy as O
    rs = np.random.RandomState(3)
    p0 = rs.standard_normal(1000).astype(np.float32)
    tp = torch.nn.Parameter(torch.from_numpy(p0.copy()))
    opt = torch.optim.SGD([tp], **kw)
    p, buf = p0.copy(), np.zeros_like(p0)
    for t in range(1, 6):
        g = (rs.standard_normal(1000) * 10.0 ** rs.uniform(-4, 1, 1000)).astype(np.float32)
        tp.grad = torch.from_numpy(g.copy())
        opt.step()
        O.sgd_update(p, g, buf, t, **kw)
        np.testing.assert_allclose(p, tp.detach().numpy(), rtol=2e-6, atol=1e-7)
