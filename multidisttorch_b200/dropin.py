"""Drop-in hooks that put the B200 engine behind the torch names the reference
scripts call directly (SURVEY §8b "compat mode").

``vae-hpo.py`` builds its trial with
``torch.nn.parallel.DistributedDataParallel(model, process_group=group)`` and
``optim.Adam(model.parameters(), lr=1e-3)`` (ref:vae-hpo.py:130-131) and then
runs ``zero_grad / model(data) / loss_function / backward / step``
(ref:vae-hpo.py:69-74).  ``install()`` (called when ``utils`` is imported on a
CUDA box) replaces those two names with factories that

* recognise the reference's VAE (five ``nn.Linear`` named fc1, fc21, fc22, fc3,
  fc4 with encoder/decoder-consistent shapes, fp32, on CUDA) and return an
  ``EngineModule`` / ``EngineAdam`` pair backed by ONE ``VAEEngine``;
* fall through to the real torch classes for anything else.

``torch.optim.SGD`` over the wrapped model's parameters becomes ``EngineSGD`` (the engine's flat SGD kernel; the
reference itself only builds Adam).

``torch.utils.data.DataLoader`` gets the same treatment for the reference's two MNIST loaders
(ref:vae-hpo.py:148-158): a device-resident loader that yields bit-identical batches without the host pipeline.

Inside an ``EngineModule`` the forward is one C-ABI call (tensor-core or fp32
GEMMs + fused reparameterisation, eps drawn from the same Philox stream
``torch.randn_like`` would use), the script's own ``loss_function`` still runs
in torch on the returned tensors, ``loss.backward()`` lands in
``mdt_backward`` + the subgroup gradient all-reduce, and ``optimizer.step()`` is
the fused flat Adam kernel (with the 1/group_size scale folded in).
There is no CPU path: on a box without CUDA nothing is patched.
"""
import os

import torch
import torch.distributed as dist
import torch.utils.data as _tud
from torch import nn

from .engine import REF_NAMES, VAEEngine

_ORIG = {}
_ENGINE_OF_PARAM = {}     # id(parameter) -> EngineModule


def _next_pow2(n):
    p = 1
    while p < n:
        p *= 2
    return p


def recognise(module):
    """(in_dim, hidden, latent) if ``module`` has the reference VAE's structure, else None."""
    names = ("fc1", "fc21", "fc22", "fc3", "fc4")
    if not all(isinstance(getattr(module, n, None), nn.Linear) for n in names):
        return None
    if len(list(module.parameters())) != 10:
        return None
    D, H, Z = module.fc1.in_features, module.fc1.out_features, module.fc21.out_features
    ok = (module.fc21.in_features == H and module.fc22.in_features == H and module.fc22.out_features == Z
          and module.fc3.in_features == Z and module.fc3.out_features == H
          and module.fc4.in_features == H and module.fc4.out_features == D)
    p = module.fc1.weight
    if not ok or p.dtype != torch.float32 or not p.is_cuda:
        return None
    if D % 4 or H % 4 or Z % 4:
        return None
    return D, H, Z


class _Mismatch(Exception):
    """The module has the reference VAE's layers but does not compute the reference VAE's forward."""


def _cuda_leg_is_nccl(pg):
    """True when collectives on CUDA tensors of ``pg`` go through NCCL.  ``dist.get_backend`` returns the backend
    STRING the group was created with -- "nccl", but also "cpu:gloo,cuda:nccl" for the groups this repo's own
    setup_ddp creates -- so the decision is taken on the backend object that serves the CUDA device."""
    try:
        backend = (pg if pg is not None else dist.group.WORLD)._get_backend(torch.device("cuda"))
        return isinstance(backend, dist.ProcessGroupNCCL)
    except Exception:
        name = str(dist.get_backend(pg))
        return name == "nccl" or "cuda:nccl" in name


def _notice(msg, pg):
    if os.environ.get("MDT_DROPIN_QUIET", "0") == "1":
        return
    if not dist.is_initialized() or pg is None or dist.get_rank(pg) == 0:
        print(f"[multidisttorch_b200] {msg}", flush=True)


class _VAEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, anchor, owner):
        ctx.owner = owner
        recon, mu, logvar = owner._engine_forward(x)
        return recon, mu, logvar

    @staticmethod
    def backward(ctx, d_recon, d_mu, d_logvar):
        ctx.owner._engine_backward(d_recon, d_mu, d_logvar)
        return None, None, None


class EngineModule(nn.Module):
    """Stands where DistributedDataParallel(model, process_group=group) stood."""

    def __init__(self, module, dims, process_group=None):
        super().__init__()
        self.module = module
        self.dims = dims
        self.process_group = process_group
        self.gemm_mode = os.environ.get("MDT_GEMM", "tf32x3")
        self.engine = None
        self._gen = torch.cuda.default_generators[module.fc1.weight.device.index]
        self._expect_offset = None
        self._build(int(os.environ.get("MDT_MAX_BATCH", "1024")), first=True)

    # -- engine lifetime ---------------------------------------------------------------------
    def _build(self, max_batch, first=False):
        D, H, Z = self.dims
        dev = self.module.fc1.weight.device
        old = self.engine
        eng = VAEEngine(in_dim=D, hidden=H, latent=Z, max_batch=max_batch, gemm_mode=self.gemm_mode,
                        use_graph=os.environ.get("MDT_GRAPH", "1") != "0", group=self.process_group, device=dev)
        if first:
            eng.load_reference_state_dict(self.module.state_dict())
            try:
                self._verify(eng)
            except _Mismatch:
                eng.close()
                raise
            if eng.group_size > 1:
                eng.broadcast_params()          # DDP constructor: parameters of group-rank 0 win
        else:
            eng.params.copy_(old.params); eng.exp_avg.copy_(old.exp_avg); eng.exp_avg_sq.copy_(old.exp_avg_sq)
            eng.grads.copy_(old.grads)
            t, _ = old.state()
            eng.params_changed()
            if getattr(self, "_sgd_h", None) is not None:
                eng.set_sgd(*self._sgd_h, completed_steps=t)
            else:
                lr, betas, eps = getattr(self, "_adam_h", (1e-3, (0.9, 0.999), 1e-8))
                eng.set_adam(lr, betas, eps, completed_steps=t)
            old.close()
        self.engine = eng
        # the script's parameters become views of the flat buffers (so decode(), state_dict(), .grad keep working)
        pv, gv = eng.named_parameter_views(), eng.named_parameter_views(eng.grads)
        for name, p in self.module.named_parameters():
            p.data = pv[name]
            p.grad = gv[name]
            _ENGINE_OF_PARAM[id(p)] = self
        self._sync_rng(force=True)

    def _verify(self, eng):
        """Behaviour check before the swap: the module's OWN forward (eps pinned by intercepting ``torch.randn_like``
        for the duration of the call) must agree with the engine on a small batch.  A model that merely has layers
        named fc1 .. fc4 but another activation, dropout, a scaled reparameterisation or a different output tuple
        raises _Mismatch and is left to torch."""
        D, H, Z = self.dims
        dev = eng.device
        g = torch.Generator(device="cpu").manual_seed(20240607)
        x = torch.rand(4, D, generator=g).to(dev)
        eps = torch.randn(4, Z, generator=g).to(dev)
        orig = torch.randn_like
        calls = []

        def pinned(like, *a, **k):
            calls.append(tuple(like.shape))
            return eps.to(like.dtype) if tuple(like.shape) == (4, Z) else orig(like, *a, **k)

        was_training = self.module.training
        try:
            torch.randn_like = pinned
            with torch.no_grad():
                out = self.module(x)
        except Exception as exc:          # the module cannot even take a [B, D] batch
            raise _Mismatch(f"forward failed on a [4, {D}] batch: {exc}")
        finally:
            torch.randn_like = orig
            self.module.train(was_training)
        if not (isinstance(out, tuple) and len(out) == 3 and all(torch.is_tensor(t) for t in out)):
            raise _Mismatch("forward does not return (recon, mu, logvar)")
        if [tuple(t.shape) for t in out] != [(4, D), (4, Z), (4, Z)] or calls != [(4, Z)]:
            raise _Mismatch("forward has other shapes / another source of eps than the reference VAE")
        got = eng.forward(x, eps)
        tol = 3e-2 if self.gemm_mode in ("tf32", "bf16") else 1e-4
        for name, a, b in zip(("recon", "mu", "logvar"), got, out):
            err = float((a - b).norm() / b.norm().clamp_min(1e-30))
            if not err <= tol:
                raise _Mismatch(f"{name} differs from the module's own forward (rel. L2 {err:.2e} > {tol:.0e})")

    def _sync_rng(self, force=False):
        off = int(self._gen.get_offset())
        if force or off != self._expect_offset:
            self.engine.set_rng(int(self._gen.initial_seed()), off)
            self._expect_offset = off

    # -- forward / backward --------------------------------------------------------------------
    def _engine_forward(self, x):
        x = x.reshape(-1, self.dims[0])
        if x.dtype != torch.float32 or not x.is_contiguous():
            x = x.contiguous().float()
        if x.shape[0] > self.engine.max_batch:
            self._build(_next_pow2(x.shape[0]))
        self._sync_rng()
        out = self.engine.forward(x, None)
        # keep torch's CUDA generator in lock-step with what randn_like would have consumed
        B, Z = x.shape[0], self.dims[2]
        if B > 0:
            props = torch.cuda.get_device_properties(x.device)
            grid = min((B * Z + 255) // 256, props.multi_processor_count * (props.max_threads_per_multi_processor // 256))
            inc = ((B * Z - 1) // (256 * grid * 4) + 1) * 4
            self._expect_offset += inc
            self._gen.set_offset(self._expect_offset)
        return out

    def _engine_backward(self, d_recon, d_mu, d_logvar):
        B = d_recon.shape[0] if d_recon is not None else (d_mu.shape[0] if d_mu is not None else 0)
        D, _, Z = self.dims
        dev = self.engine.device
        if d_recon is None:
            d_recon = torch.zeros(B, D, device=dev)
        if d_mu is None:
            d_mu = torch.zeros(B, Z, device=dev)
        if d_logvar is None:
            d_logvar = torch.zeros(B, Z, device=dev)
        self.engine.backward(d_recon, d_mu, d_logvar)
        self.engine.allreduce_grads()        # SUM over the trial group; Adam applies 1/group_size

    def forward(self, x):
        if torch.is_grad_enabled() and self.training:
            return _VAEFn.apply(x, self.module.fc1.weight, self)
        return self._engine_forward(x)


class EngineAdam:
    """Stands where optim.Adam(model.parameters(), lr=1e-3) stood."""

    def __init__(self, owner, lr, betas, eps):
        self.owner = owner
        from . import hpo
        if hpo.overridden():           # per-trial hyper-parameters from MDT_HPO_* (SURVEY 8f-4), keyed by the trial's group
            cfg = hpo.trial_config(hpo.group_index(owner.process_group), lr=lr, beta1=betas[0], beta2=betas[1], eps=eps)
            lr, betas, eps = cfg.lr, cfg.betas, cfg.eps
        self.defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=0, amsgrad=False)
        self.param_groups = [dict(self.defaults, params=list(owner.module.parameters()))]
        owner._adam_h, owner._sgd_h = (lr, betas, eps), None
        owner.engine.set_adam(lr, betas, eps, 0)

    def zero_grad(self, set_to_none=True):
        pass        # gradients are overwritten by every backward, never accumulated

    def step(self, closure=None):
        self.owner.engine.adam_step()
        return None

    def state_dict(self):
        t, _ = self.owner.engine.state()
        eng = self.owner.engine
        return {"step": t, "exp_avg": eng.exp_avg.clone(), "exp_avg_sq": eng.exp_avg_sq.clone(),
                "param_groups": [{k: v for k, v in self.param_groups[0].items() if k != "params"}]}


class EngineSGD:
    """Stands where ``optim.SGD(model.parameters(), lr, momentum, dampening, weight_decay, nesterov)`` would stand
    (BASELINE north_star (c); the reference itself instantiates Adam, ref:vae-hpo.py:131).  ``step()`` is the flat SGD
    kernel (mdt_engine_set_sgd + mdt_adam_step); the engine's exp_avg buffer holds the momentum buffer."""

    def __init__(self, owner, lr, momentum, dampening, weight_decay, nesterov):
        self.owner = owner
        from . import hpo
        if hpo.overridden():           # MDT_HPO_LR applies to whichever optimizer the script builds
            lr = hpo.trial_config(hpo.group_index(owner.process_group), lr=lr).lr
        self.defaults = dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay, nesterov=nesterov)
        self.param_groups = [dict(self.defaults, params=list(owner.module.parameters()))]
        owner._sgd_h = (lr, momentum, dampening, weight_decay, nesterov)
        owner.engine.exp_avg.zero_()   # torch creates the momentum buffer on the first step
        owner.engine.set_sgd(lr, momentum, dampening, weight_decay, nesterov, completed_steps=0)

    def zero_grad(self, set_to_none=True):
        pass

    def step(self, closure=None):
        self.owner.engine.adam_step()      # the engine's optimizer step: SGD once set_sgd was called
        return None

    def state_dict(self):
        t, _ = self.owner.engine.state()
        return {"step": t, "momentum_buffer": self.owner.engine.exp_avg.clone(),
                "param_groups": [{k: v for k, v in self.param_groups[0].items() if k != "params"}]}


def _engine_owner(plist):
    """The EngineModule whose ten parameters `plist` is, or None."""
    if not plist or isinstance(plist[0], dict):
        return None
    owners = {id(_ENGINE_OF_PARAM.get(id(p))) for p in plist}
    owner = _ENGINE_OF_PARAM.get(id(plist[0]))
    return owner if owner is not None and len(owners) == 1 and len(plist) == 10 else None


def _make_sgd_shim(orig_sgd):
    class SGD(orig_sgd):
        def __new__(cls, params=None, lr=1e-3, momentum=0, dampening=0, weight_decay=0, nesterov=False, **kwargs):
            plist = list(params) if params is not None else []
            if cls is SGD:
                owner = _engine_owner(plist)
                plain = (owner is not None and not kwargs and not isinstance(lr, torch.Tensor)
                         and not (nesterov and (momentum <= 0 or dampening != 0)))     # torch raises for that one itself
                if plain:
                    return EngineSGD(owner, float(lr), float(momentum), float(dampening), float(weight_decay), bool(nesterov))
            obj = super().__new__(cls)
            obj._mdt_params = plist          # a generator argument was consumed above: hand the list to __init__
            return obj

        def __init__(self, params=None, *args, **kwargs):
            super().__init__(self.__dict__.pop("_mdt_params", params), *args, **kwargs)

    SGD.__qualname__ = SGD.__name__ = "SGD"
    return SGD


class DeviceLoader:
    """Stands where ``DataLoader(datasets.MNIST(..., transform=ToTensor()), batch_size=B, shuffle=False,
    sampler=DistributedSampler(...))`` stood (ref:vae-hpo.py:135-158).

    The reference's loader decodes every sample through PIL and converts uint8 -> float on the host, ~59 k samples/s per
    process -- 25-40x below what the step consumes.  Here the uint8 image table is uploaded once; each batch is gathered
    and scaled (/255, bit-identical to ToTensor()) on the device in the sampler's own index order, and handed to the
    script as a CUDA tensor [B, 1, 28, 28] (its ``data.to(device)`` is then a no-op).  Labels follow as a CUDA tensor.
    ``len()``, ``.dataset``, ``.batch_size`` and ``.sampler`` behave like the DataLoader's."""

    def __init__(self, dataset, batch_size, sampler, drop_last):
        from .data import DeviceDataset
        self.dataset, self.batch_size, self.sampler, self.drop_last = dataset, int(batch_size), sampler, bool(drop_last)
        self._dd = DeviceDataset(dataset.data)
        self._targets = torch.as_tensor(dataset.targets).to(self._dd.device)
        self._shape = tuple(dataset.data.shape[1:])

    def _n(self):
        return len(self.sampler) if self.sampler is not None else len(self.dataset)

    def __len__(self):
        n, b = self._n(), self.batch_size
        return n // b if self.drop_last else (n + b - 1) // b

    def __iter__(self):
        order = list(iter(self.sampler)) if self.sampler is not None else range(len(self.dataset))
        idx = torch.as_tensor(list(order), dtype=torch.int64).to(self._dd.device)
        b = self.batch_size
        for i in range(0, idx.numel(), b):
            sel = idx[i:i + b]
            if sel.numel() < b and self.drop_last:
                return
            x = torch.empty((sel.numel(), 1) + self._shape, dtype=torch.float32, device=self._dd.device)
            self._dd.gather(sel, x)
            yield x, self._targets[sel]


def _wants_device_loader(dataset, args, kwargs):
    """The reference's two loaders and nothing else: torchvision MNIST with exactly ToTensor() as transform, default
    collation, no workers, sequential or DistributedSampler order."""
    if args or not set(kwargs) <= {"batch_size", "shuffle", "sampler", "drop_last"} or kwargs.get("shuffle"):
        return False
    try:
        from torchvision import datasets, transforms
    except Exception:
        return False
    if type(dataset) is not datasets.MNIST or type(dataset.transform) is not transforms.ToTensor:
        return False
    if dataset.target_transform is not None or getattr(dataset, "transforms", None) is None:
        return False
    smp = kwargs.get("sampler")
    from torch.utils.data.distributed import DistributedSampler
    if smp is not None and type(smp) is not DistributedSampler:
        return False
    return torch.cuda.is_available() and os.environ.get("MDT_DEVICE_LOADER", "1") != "0"


def _wants_engine_ddp(module, args, kwargs):
    dims = recognise(module) if not args and set(kwargs) <= {"process_group"} else None
    if dims is None:
        return None
    pg = kwargs.get("process_group")
    if dist.is_initialized() and dist.get_world_size(pg) > 1 and not _cuda_leg_is_nccl(pg):
        return None         # e.g. DDP_BACKEND=gloo with CUDA tensors: leave to torch
    return dims


def _make_shims(orig_ddp, orig_adam):
    """Subclasses of the real torch classes whose constructors hand back the engine-backed
    objects for the reference VAE and behave exactly like the originals otherwise
    (isinstance checks and further subclassing keep working)."""

    class DistributedDataParallel(orig_ddp):
        def __new__(cls, module=None, *args, **kwargs):
            if cls is DistributedDataParallel and isinstance(module, nn.Module):
                dims = _wants_engine_ddp(module, args, kwargs)
                if dims is not None:
                    pg = kwargs.get("process_group")
                    try:
                        em = EngineModule(module, dims, pg)
                        _notice(f"DistributedDataParallel(VAE {dims[0]}-{dims[1]}-{dims[2]}) -> B200 engine "
                                f"(gemm_mode={em.gemm_mode}, group of {em.engine.group_size}, "
                                f"gradient exchange: {'peer memory' if em.engine.p2p else 'NCCL' if em.engine.group_size > 1 else 'none'})", pg)
                        return em
                    except _Mismatch as why:
                        _notice(f"module has the reference VAE's layers but {why}: left to torch DDP", pg)
            return super().__new__(cls)

    class Adam(orig_adam):
        def __new__(cls, params=None, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False, **kwargs):
            plist = list(params) if params is not None else []
            if cls is Adam and plist and not isinstance(plist[0], dict):
                owners = {id(_ENGINE_OF_PARAM.get(id(p))) for p in plist}
                owner = _ENGINE_OF_PARAM.get(id(plist[0]))
                plain = (owner is not None and len(owners) == 1 and len(plist) == 10 and weight_decay == 0
                         and not amsgrad and not kwargs and not isinstance(lr, torch.Tensor))
                if plain:
                    return EngineAdam(owner, lr, betas, eps)
            obj = super().__new__(cls)
            obj._mdt_params = plist          # a generator argument was consumed above: hand the list to __init__
            return obj

        def __init__(self, params=None, *args, **kwargs):
            super().__init__(self.__dict__.pop("_mdt_params", params), *args, **kwargs)

    DistributedDataParallel.__qualname__ = DistributedDataParallel.__name__ = "DistributedDataParallel"
    Adam.__qualname__ = Adam.__name__ = "Adam"
    return DistributedDataParallel, Adam


def _make_loader_shim(orig_loader):
    class DataLoader(orig_loader):
        def __new__(cls, dataset=None, *args, **kwargs):
            if cls is DataLoader and dataset is not None and _wants_device_loader(dataset, args, kwargs):
                from . import hpo
                batch = kwargs.get("batch_size", 1)
                smp = kwargs.get("sampler")
                if smp is not None and hpo.overridden():     # MDT_HPO_BATCH: the trial's own batch size (train loader only;
                    batch = hpo.trial_config(smp.rank, batch_size=batch).batch_size   # the sampler's rank is the group id)
                _notice(f"DataLoader(MNIST, batch_size={batch}) -> device-resident loader "
                        f"({len(dataset)} images in HBM, gather + /255 on the GPU)", None)
                return DeviceLoader(dataset, batch, smp, kwargs.get("drop_last", False))
            return super().__new__(cls)

    DataLoader.__qualname__ = DataLoader.__name__ = "DataLoader"
    return DataLoader


def install():
    """Patch torch.nn.parallel.DistributedDataParallel, torch.optim.Adam / SGD and torch.utils.data.DataLoader
    (idempotent)."""
    if _ORIG or not torch.cuda.is_available():
        return False
    _ORIG["ddp"] = torch.nn.parallel.DistributedDataParallel
    _ORIG["adam"] = torch.optim.Adam
    ddp, adam = _make_shims(_ORIG["ddp"], _ORIG["adam"])
    torch.nn.parallel.DistributedDataParallel = ddp
    torch.optim.Adam = adam
    _ORIG["sgd"] = torch.optim.SGD
    torch.optim.SGD = _make_sgd_shim(_ORIG["sgd"])
    _ORIG["loader"] = _tud.DataLoader
    _tud.DataLoader = _make_loader_shim(_ORIG["loader"])
    return True


def uninstall():
    if not _ORIG:
        return
    torch.nn.parallel.DistributedDataParallel = _ORIG.pop("ddp")
    torch.optim.Adam = _ORIG.pop("adam")
    if "sgd" in _ORIG:
        torch.optim.SGD = _ORIG.pop("sgd")
    if "loader" in _ORIG:
        _tud.DataLoader = _ORIG.pop("loader")
