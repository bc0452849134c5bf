#!/usr/bin/env python
"""Generate tests/golden/* by running the REAL reference (TEST INFRASTRUCTURE).

Run in the build container only (needs /root/reference, which does not exist
on the GPU box):

    python oracle/make_golden.py            # rewrites tests/golden/*.npz, api.json

The reference module is loaded from /root/reference/vae-hpo.py with importlib
(hyphenated file name) and driven exactly as ref:vae-hpo.py:69-74 drives it:
``zero_grad -> model(x) -> loss_function -> backward -> step``.  Only two
things are injected, because the reference seeds nothing:

* initial parameters: ``oracle.vae_numpy.init_params(seed)`` copied into the
  reference ``VAE`` through ``load_state_dict``;
* eps: ``torch.randn_like`` is patched to return the next supplied eps tensor.

Inputs (params0, x, eps) are regenerated from numpy seeds by the tests, so the
fixtures hold outputs only: per-step losses, step-0 activations and gradients,
parameters / Adam moments after N steps -- large tensors as fingerprints
(float64 sum, l2 and every 61st element), small ones in full.

While generating, the script also asserts that ``oracle.vae_numpy`` agrees
with the reference on the FULL tensors (not just the fingerprints).
"""
from __future__ import annotations

import argparse
import importlib.util
import io
import json
import os
import subprocess
import sys
import tempfile
from contextlib import redirect_stdout

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = "/root/reference"
GOLD = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, ROOT)

from oracle import vae_numpy as O  # noqa: E402

# fixed seeds shared with tests/ (tests/golden_cases.py re-exports them)
PARAM_SEED = 1234
DATA_SEED = 4321

# our flat segment -> reference state_dict construction
def ref_state_from_oracle(p, torch):
    Z = p["W2x"].shape[0] // 2
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a))
    return {
        "fc1.weight": t(p["W1"]), "fc1.bias": t(p["b1"]),
        "fc21.weight": t(p["W2x"][:Z]), "fc21.bias": t(p["b2x"][:Z]),
        "fc22.weight": t(p["W2x"][Z:]), "fc22.bias": t(p["b2x"][Z:]),
        "fc3.weight": t(p["W3"]), "fc3.bias": t(p["b3"]),
        "fc4.weight": t(p["W4"]), "fc4.bias": t(p["b4"]),
    }


def oracle_from_ref(named, attr=None):
    """dict name->tensor (reference naming) -> oracle segment dict."""
    g = lambda k: (named[k] if attr is None else attr(named[k])).detach().cpu().numpy()
    return {
        "W1": g("fc1.weight"), "b1": g("fc1.bias"),
        "W2x": np.concatenate([g("fc21.weight"), g("fc22.weight")], 0),
        "b2x": np.concatenate([g("fc21.bias"), g("fc22.bias")], 0),
        "W3": g("fc3.weight"), "b3": g("fc3.bias"),
        "W4": g("fc4.weight"), "b4": g("fc4.bias"),
    }


def load_reference():
    import torch
    if REF not in sys.path:
        sys.path.insert(0, REF)  # so `from utils import *` resolves to the reference's utils.py
    spec = importlib.util.spec_from_file_location("ref_vae_hpo", os.path.join(REF, "vae-hpo.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return torch, mod


class EpsInjector:
    """Patch torch.randn_like so VAE.reparameterize (ref:vae-hpo.py:35) draws
    the eps we supply."""

    def __init__(self, torch):
        self.torch = torch
        self.queue = []

    def __enter__(self):
        self._orig = self.torch.randn_like
        self.torch.randn_like = lambda like, *a, **k: self.queue.pop(0).to(like.dtype)
        return self

    def __exit__(self, *exc):
        self.torch.randn_like = self._orig


def put_fp(out, prefix, seg_dict, stride=61):
    for k, v in seg_dict.items():
        fp = O.fingerprint(v, stride)
        out[f"{prefix}.{k}.sum"] = fp["sum"]
        out[f"{prefix}.{k}.l2"] = fp["l2"]
        out[f"{prefix}.{k}.sample"] = fp["sample"]


def rel_l2(a, b):
    a = np.asarray(a, np.float64).ravel()
    b = np.asarray(b, np.float64).ravel()
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def batch_seed(step, rank=0):
    return DATA_SEED + 1000 * rank + step


def widen(ref_model, torch, H):
    """BASELINE config 5 ("widened hidden dim 4096"): the reference's VAE with its five nn.Linear
    re-sized to hidden width H.  forward / loss / autograd stay the reference's own code."""
    if H == 400:
        return ref_model
    nn = torch.nn
    ref_model.fc1, ref_model.fc21, ref_model.fc22 = nn.Linear(784, H), nn.Linear(H, 20), nn.Linear(H, 20)
    ref_model.fc3, ref_model.fc4 = nn.Linear(20, H), nn.Linear(H, 784)
    return ref_model


def run_single(B, n_steps, tag, H=400, stride=61):
    """One rank, no DDP: ref:vae-hpo.py:69-74 verbatim."""
    torch, ref = load_reference()
    torch.set_num_threads(8)
    dims = O.Dims(H=H)
    p0 = O.init_params(PARAM_SEED, dims)
    model = widen(ref.VAE(), torch, H)
    model.load_state_dict(ref_state_from_oracle(p0, torch))
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    orc = O.Trainer(p0, dims)
    out = {"B": B, "n_steps": n_steps, "H": H, "stride": stride}
    losses = []
    _put = put_fp
    put_fp_l = lambda o, pre, d: _put(o, pre, d, stride)
    with EpsInjector(torch) as inj:
        for t in range(n_steps):
            x, eps = O.make_batch(batch_seed(t), B, dims)
            data = torch.from_numpy(x).view(B, 1, 28, 28)
            inj.queue.append(torch.from_numpy(eps))
            opt.zero_grad()
            recon, mu, logvar = model(data)
            loss = ref.loss_function(recon, data, mu, logvar)
            loss.backward()
            losses.append(loss.item())
            if t == 0:
                put_fp_l(out, "step0", {"recon": recon.detach().numpy(), "mu": mu.detach().numpy(),
                                      "logvar": logvar.detach().numpy()})
                g_ref = oracle_from_ref(dict(model.named_parameters()), attr=lambda q: q.grad)
                put_fp_l(out, "grad0", g_ref)
            opt.step()
            # oracle in lock-step, full-tensor check
            lo, go = orc.step([(x, eps)])
            assert abs(lo[0] - losses[-1]) <= 2e-6 * abs(losses[-1]), (tag, t, lo[0], losses[-1])
            if t == 0:
                for k in g_ref:
                    assert rel_l2(go[k], g_ref[k]) < 2e-5, (tag, "grad", k, rel_l2(go[k], g_ref[k]))
    pN = oracle_from_ref(dict(model.named_parameters()))
    for k in pN:
        # H=4096: K=4096 dot products differ between numpy's and torch's fp32 summation order by ~1e-6, and the
        # first Adam steps (update ~ lr * sign(g)) amplify that ~50x (measured 6.9e-5 on W1 after 3 steps)
        assert rel_l2(orc.p[k], pN[k]) < (1e-5 if H == 400 else 2e-4), (tag, "param", k, rel_l2(orc.p[k], pN[k]))
    put_fp_l(out, "paramN", pN)
    st = opt.state_dict()["state"]
    names = [n for n, _ in model.named_parameters()]
    mN = oracle_from_ref({n: st[i]["exp_avg"] for i, n in enumerate(names)})
    vN = oracle_from_ref({n: st[i]["exp_avg_sq"] for i, n in enumerate(names)})
    put_fp_l(out, "mN", mN)
    put_fp_l(out, "vN", vN)
    out["losses"] = np.array(losses, dtype=np.float64)
    np.savez_compressed(os.path.join(GOLD, f"{tag}.npz"), **out)
    print(f"[golden] {tag}: losses {losses[0]:.4f} .. {losses[-1]:.4f}")


def run_eval(B, tag):
    """Forward-only eval path, ref:vae-hpo.py:101-105."""
    torch, ref = load_reference()
    dims = O.Dims()
    p0 = O.init_params(PARAM_SEED, dims)
    model = ref.VAE()
    model.load_state_dict(ref_state_from_oracle(p0, torch))
    model.eval()
    x, eps = O.make_batch(batch_seed(77), B, dims)
    with torch.no_grad(), EpsInjector(torch) as inj:
        inj.queue.append(torch.from_numpy(eps))
        data = torch.from_numpy(x).view(B, 1, 28, 28)
        recon, mu, logvar = model(data)
        loss = ref.loss_function(recon, data, mu, logvar).item()
        zs = np.random.RandomState(99).standard_normal((64, dims.Z)).astype(np.float32)
        sample = model.decode(torch.from_numpy(zs)).numpy()   # ref:vae-hpo.py:164-165
    act = O.forward(p0, x, eps, dims)
    assert rel_l2(act["recon"], recon.numpy()) < 1e-5
    out = {"B": B, "loss": np.float64(loss)}
    put_fp(out, "eval", {"recon": recon.numpy(), "mu": mu.numpy(), "logvar": logvar.numpy(), "decode64": sample})
    np.savez_compressed(os.path.join(GOLD, f"{tag}.npz"), **out)
    print(f"[golden] {tag}: loss {loss:.4f}")


def run_loss_edge(tag):
    """Known answers for loss_function and its autograd at the clamp/saturation
    corners (ref:vae-hpo.py:49-58)."""
    torch, ref = load_reference()
    rs = np.random.RandomState(7)
    B, D, Z = 8, 784, 20
    recon = rs.random_sample((B, D)).astype(np.float32)
    x = rs.random_sample((B, D)).astype(np.float32)
    specials = np.array([0.0, 1.0, 1e-45, 1e-38, 1e-20, 1e-12, 1 - 2.0 ** -24, 1 - 2.0 ** -23, 0.5,
                         2.0 ** -24, 1e-7, 0.999999], dtype=np.float32)
    recon[0, :len(specials)] = specials
    recon[1, :len(specials)] = specials
    x[0, :len(specials)] = 0.0
    x[1, :len(specials)] = 1.0
    recon[2, :len(specials)] = specials
    x[2, :len(specials)] = 0.25
    mu = (rs.standard_normal((B, Z)) * 3).astype(np.float32)
    lv = (rs.standard_normal((B, Z)) * 4).astype(np.float32)
    lv[0, :4] = [-30.0, 20.0, 0.0, 80.0]
    tr = torch.from_numpy(recon).requires_grad_()
    tm = torch.from_numpy(mu).requires_grad_()
    tl = torch.from_numpy(lv).requires_grad_()
    L = ref.loss_function(tr, torch.from_numpy(x).view(B, 1, 28, 28), tm, tl)
    L.backward()
    b, k = O.loss_terms(recon, x, mu, lv)
    dr, dm, dl = O.loss_grads(recon, x, mu, lv)
    assert abs((b + k) - L.item()) <= 1e-6 * abs(L.item()), (b + k, L.item())
    assert np.array_equal(dr, tr.grad.numpy()) or rel_l2(dr, tr.grad.numpy()) < 1e-6
    assert rel_l2(dl, tl.grad.numpy()) < 1e-6
    np.savez_compressed(os.path.join(GOLD, f"{tag}.npz"), recon=recon, x=x, mu=mu, logvar=lv,
                        loss=np.float64(L.item()), d_recon=tr.grad.numpy(), d_mu=tm.grad.numpy(),
                        d_logvar=tl.grad.numpy())
    print(f"[golden] {tag}: loss {L.item():.6e}")


EDGE_LOGITS = np.array([-120.0, -104.0, -103.0, -100.0, -95.0, -90.0, -88.0, -87.0, -60.0, -30.0, -17.0, -1e-3, 0.0, 8.0, 15.0,
                        16.0, 16.6, 16.7, 17.0, 17.5, 20.0, 50.0, 90.0, 120.0], dtype=np.float32)


def edge_setup(p0, x):
    """fc4 rows 0..n-1 get zero weights and a bias equal to a corner logit, so recon[:, j] = sigmoid(EDGE_LOGITS[j])
    for every sample; samples 0/1/2 carry targets 0 / 1 / 0.25 there.  Shared with tests/test_gpu_parity.py."""
    n = len(EDGE_LOGITS)
    p0["W4"][:n, :] = 0.0
    p0["b4"][:n] = EDGE_LOGITS
    x[0, :n], x[1, :n], x[2, :n] = 0.0, 1.0, 0.25
    return p0, x


def run_step_edge(tag, B=8):
    """One full training step of the reference whose fc4 logits sit on the BCE clamp / sigmoid saturation
    corners (ref:vae-hpo.py:40, 50): the known answer for the step's FUSED sigmoid+BCE epilogue."""
    torch, ref = load_reference()
    torch.set_num_threads(8)
    dims = O.Dims()
    x, eps = O.make_batch(batch_seed(0), B, dims)
    p0, x = edge_setup(O.init_params(PARAM_SEED, dims), x)
    model = ref.VAE()
    model.load_state_dict(ref_state_from_oracle(p0, torch))
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    out = {"B": B, "n_steps": 1}
    with EpsInjector(torch) as inj:
        data = torch.from_numpy(x).view(B, 1, 28, 28)
        inj.queue.append(torch.from_numpy(eps))
        opt.zero_grad()
        recon, mu, logvar = model(data)
        recon.retain_grad()
        loss = ref.loss_function(recon, data, mu, logvar)
        loss.backward()
        n = len(EDGE_LOGITS)
        out["recon_edge"] = recon.detach().numpy()[:3, :n].copy()
        out["d_recon_edge"] = recon.grad.numpy()[:3, :n].copy()
        g_ref = oracle_from_ref(dict(model.named_parameters()), attr=lambda q: q.grad)
        put_fp(out, "grad0", g_ref)
        out["grad0.b4.edge"] = g_ref["b4"][:n].copy()
        opt.step()
    lo, go = O.Trainer(p0, dims).step([(x, eps)])
    assert abs(lo[0] - loss.item()) <= 2e-6 * abs(loss.item()), (lo[0], loss.item())
    for k in g_ref:
        assert rel_l2(go[k], g_ref[k]) < 2e-5, (tag, k, rel_l2(go[k], g_ref[k]))
    put_fp(out, "paramN", oracle_from_ref(dict(model.named_parameters())))
    out["losses"] = np.array([loss.item()], dtype=np.float64)
    np.savez_compressed(os.path.join(GOLD, f"{tag}.npz"), **out)
    print(f"[golden] {tag}: loss {loss.item():.4f}; recon at the corners", out["recon_edge"][0])


# ----------------------------------------------------------------------------
# multi-process legs (real reference utils + real torch DDP on gloo)
# ----------------------------------------------------------------------------

def worker_ddp(args):
    """Runs inside one rank: reference setup_ddp/setup_ddp_groups + DDP-wrapped
    reference VAE (ref:vae-hpo.py:129-131), n_steps of ref:vae-hpo.py:69-74."""
    torch, ref = load_reference()
    import torch.distributed as dist
    torch.set_num_threads(2)
    buf = io.StringIO()
    with redirect_stdout(buf):
        ref.setup_ddp()
        groups = ref.setup_ddp_groups(args.ngroups)
    wr = dist.get_rank()
    membership = [int(dist.get_rank(g)) for g in groups]
    res = {"membership": np.array(membership), "stdout": buf.getvalue()}
    gid = [i for i, m in enumerate(membership) if m >= 0]
    if args.steps > 0 and gid:
        gid = gid[0]
        group = groups[gid]
        grank = dist.get_rank(group)
        dims = O.Dims()
        # group-rank 0 holds the real init; others start different to prove the ctor broadcast
        p0 = O.init_params(PARAM_SEED + gid if grank == 0 else 999 + wr, dims)
        model = ref.VAE()
        model.load_state_dict(ref_state_from_oracle(p0, torch))
        model = torch.nn.parallel.DistributedDataParallel(model, process_group=group)
        opt = torch.optim.Adam(model.parameters(), lr=1e-3)
        losses = []
        with EpsInjector(torch) as inj:
            for t in range(args.steps):
                x, eps = O.make_batch(batch_seed(t, wr) + args.seed_shift, args.batch, dims)
                data = torch.from_numpy(x).view(args.batch, 1, 28, 28)
                inj.queue.append(torch.from_numpy(eps))
                opt.zero_grad()
                recon, mu, logvar = model(data)
                loss = ref.loss_function(recon, data, mu, logvar)
                loss.backward()
                losses.append(loss.item())
                opt.step()
        res["losses"] = np.array(losses, dtype=np.float64)
        res["gid"] = gid
        pb = io.StringIO()
        with redirect_stdout(pb):
            ref.print0("hello", 3, process_group=group)
        res["print0"] = pb.getvalue()
        pN = oracle_from_ref(dict(model.module.named_parameters()))
        for k, v in pN.items():
            res[f"paramN.{k}"] = v
    if args.steps > 0:
        dist.barrier()
    np.savez(os.path.join(args.tmp, f"rank{wr}.npz"), **res)
    dist.destroy_process_group()


def spawn(world, ngroups, steps, batch, port, seed_shift=0):
    tmp = tempfile.mkdtemp(prefix="mdt_golden_")
    procs = []
    for r in range(world):
        env = dict(os.environ)
        env.update(OMPI_COMM_WORLD_SIZE=str(world), OMPI_COMM_WORLD_RANK=str(r), MASTER_PORT=str(port),
                   MASTER_ADDR="127.0.0.1", DDP_BACKEND="gloo", OMP_NUM_THREADS="2")
        procs.append(subprocess.Popen([sys.executable, os.path.abspath(__file__), "--worker", "--ngroups", str(ngroups),
                                       "--steps", str(steps), "--batch", str(batch), "--tmp", tmp,
                                       "--seed-shift", str(seed_shift)], env=env))
    for p in procs:
        assert p.wait(timeout=600) == 0
    return [dict(np.load(os.path.join(tmp, f"rank{r}.npz"), allow_pickle=False)) for r in range(world)]


def run_ddp(world, ngroups, steps, batch, tag, port, seed_shift=0):
    """seed_shift moves the data seeds of a case.  ReLU is discontinuous: a pre-activation within fp32 rounding of zero
    has a mask that depends on the summation order, and one flipped unit moves a row of W1 / W3 by O(lr) within a few
    Adam steps (the first ddp_2x2_b16 fixture had |a1| = 3.4e-8 and the GPU differed from it by 2.9e-3 on W1 while its
    losses agreed to 2e-7).  Small cases are therefore generated from seeds whose smallest |pre-activation| (fc1 and
    fc3, every rank, every step; recorded as `relu_margin`) is >= 1.5e-6; for the 3072-sample 2x4 case no such seed
    exists statistically, and its consumers use the looser parameter bound when `relu_margin` is below that."""
    ranks = spawn(world, ngroups, steps, batch, port, seed_shift)
    out = {"world": world, "ngroups": ngroups, "B": batch, "n_steps": steps, "seed_shift": seed_shift}
    margin = 1.0
    k = world // ngroups
    for gid in range(ngroups):
        members = list(range(k * gid, k * gid + k))
        # every rank of a group must end with identical parameters
        for r in members[1:]:
            for name in O.SEGMENTS:
                assert np.array_equal(ranks[r][f"paramN.{name}"], ranks[members[0]][f"paramN.{name}"]), (gid, r, name)
        put_fp(out, f"g{gid}.paramN", {n: ranks[members[0]][f"paramN.{n}"] for n in O.SEGMENTS})
        for r in members:
            out[f"losses.rank{r}"] = ranks[r]["losses"]
            out[f"print0.rank{r}"] = np.array(str(ranks[r]["print0"]))
        # oracle check (full tensors)
        orc = O.Trainer(O.init_params(PARAM_SEED + gid), O.Dims())
        for t in range(steps):
            for r in members:
                x_, eps_ = O.make_batch(batch_seed(t, r) + seed_shift, batch)
                act_ = O.forward(orc.p, x_, eps_)
                a1_ = x_ @ orc.p["W1"].T + orc.p["b1"]
                a3_ = act_["z"] @ orc.p["W3"].T + orc.p["b3"]
                margin = min(margin, float(np.abs(a1_).min()), float(np.abs(a3_).min()))
            lo, _ = orc.step([O.make_batch(batch_seed(t, r) + seed_shift, batch) for r in members])
            for i, r in enumerate(members):
                ref_l = float(ranks[r]["losses"][t])
                assert abs(lo[i] - ref_l) <= 2e-6 * abs(ref_l), (tag, gid, t, r, lo[i], ref_l)
        for name in O.SEGMENTS:
            e = rel_l2(orc.p[name], ranks[members[0]][f"paramN.{name}"])
            assert e < 1e-5, (tag, gid, name, e)
    out["relu_margin"] = np.float64(margin)
    np.savez_compressed(os.path.join(GOLD, f"{tag}.npz"), **out)
    print(f"[golden] {tag}: ok (smallest |pre-activation| {margin:.2e})")


def run_api(port):
    """Known answers of the kept utils API, produced by the reference itself."""
    sys.path.insert(0, REF)
    import importlib
    ru = importlib.import_module("utils")
    assert ru.__file__.startswith(REF)
    api = {"parse_slurm_nodelist": {}, "init_comm_size_and_rank": [], "groups": [], "star_names": None}
    for s in ["or-condo-g04", "or-condo-g[05,07-08,13]", "or-condo-g[05,07-08,13],or-condo-h[01,12]",
              "n[098-101]", "a1,b2", "node[1-3]", "x[9-11]", "gpu-[001-003,010]", "single"]:
        api["parse_slurm_nodelist"][s] = ru.parse_slurm_nodelist(s)
    api["find_ifname_loopback"] = ru.find_ifname("127.0.0.1")
    envs = [
        {}, {"SLURM_NPROCS": "4", "SLURM_PROCID": "3"},
        {"OMPI_COMM_WORLD_SIZE": "8", "OMPI_COMM_WORLD_RANK": "5", "SLURM_NPROCS": "4", "SLURM_PROCID": "3"},
        {"RANK": "3", "WORLD_SIZE": "4"}, {"OMPI_COMM_WORLD_SIZE": "2"}, {"SLURM_NPROCS": "2", "SLURM_PROCID": "0"},
    ]
    keys = ["OMPI_COMM_WORLD_SIZE", "OMPI_COMM_WORLD_RANK", "SLURM_NPROCS", "SLURM_PROCID", "RANK", "WORLD_SIZE"]
    saved = {k: os.environ.pop(k, None) for k in keys}
    for e in envs:
        for k in keys:
            os.environ.pop(k, None)
        os.environ.update(e)
        api["init_comm_size_and_rank"].append({"env": e, "out": list(ru.init_comm_size_and_rank())})
    for k in keys:
        os.environ.pop(k, None)
    for k, v in saved.items():
        if v is not None:
            os.environ[k] = v
    ns = {}
    exec("from utils import *", ns)
    api["star_names"] = sorted(k for k in ns if not k.startswith("__"))
    for i, (w, g) in enumerate([(2, 2), (4, 2), (4, 1), (4, 4), (5, 2), (8, 3), (8, 4), (8, 2)]):
        ranks = spawn(w, g, 0, 0, port + i)
        api["groups"].append({"world": w, "ngroups": g,
                              "membership": [r["membership"].tolist() for r in ranks],
                              "stdout_rank0": str(ranks[0]["stdout"]),
                              "stdout_last": str(ranks[-1]["stdout"])})
        print(f"[golden] groups ({w},{g}) ok")
    with open(os.path.join(GOLD, "api.json"), "w") as f:
        json.dump(api, f, indent=1, sort_keys=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--worker", action="store_true")
    ap.add_argument("--ngroups", type=int, default=1)
    ap.add_argument("--steps", type=int, default=0)
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--tmp", default="")
    ap.add_argument("--seed-shift", type=int, default=0)
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    if args.worker:
        return worker_ddp(args)
    os.makedirs(GOLD, exist_ok=True)
    todo = args.only.split(",") if args.only else ["single", "eval", "edge", "ddp", "api"]
    if "single" in todo:
        run_single(128, 10, "step_b128")
        run_single(48, 3, "step_b48")
        run_single(1, 2, "step_b1")
        run_single(512, 3, "step_b512")                          # BASELINE config 4's group batch on one rank
        run_single(128, 3, "step_h4096_b128", H=4096, stride=997)   # BASELINE config 5's width
    if "eval" in todo:
        run_eval(64, "eval_b64")
    if "edge" in todo:
        run_loss_edge("loss_edge")
        run_step_edge("step_edge_b8")
    if "ddp" in todo:
        run_ddp(2, 1, 5, 32, "ddp_1x2_b32", 29610)
        run_ddp(4, 2, 3, 16, "ddp_2x2_b16", 29620, seed_shift=300000)
        run_ddp(4, 1, 3, 32, "ddp_1x4_b32", 29650, seed_shift=1500000)
        run_ddp(8, 1, 3, 16, "ddp_1x8_b16", 29660, seed_shift=200000)
        run_ddp(8, 2, 3, 128, "ddp_2x4_b128", 29670)             # BASELINE config 4: 2 trials x 4 ranks, 512/group
    if "api" in todo:
        run_api(29640)
    # print0 format known answer (from the ddp leg)
    print("[golden] done ->", GOLD)


if __name__ == "__main__":
    main()
