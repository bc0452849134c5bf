"""Python handle on one trial's B200 engine (host side of the C-ABI).

Mirrors what ref:vae-hpo.py:129-131 builds -- ``VAE().to(device)`` wrapped in
``DistributedDataParallel(process_group=group)`` plus ``optim.Adam(lr=1e-3)`` --
as ONE object that owns four flat fp32 device buffers (params, grads, exp_avg,
exp_avg_sq) and drives the hand-written sm_100a kernels through ctypes.  torch
is used for device memory, streams and torch.distributed only."""
import ctypes as C
import os

import numpy as np
import torch
import torch.distributed as dist

from . import _lib
from ._lib import MDT_EXCHANGE, MDT_GEMM, SEGMENTS, check, mdt_adam, mdt_batch_source, mdt_cfg, mdt_sgd

# reference parameter names, ref:vae-hpo.py:23-27
REF_NAMES = ("fc1.weight", "fc1.bias", "fc21.weight", "fc21.bias", "fc22.weight", "fc22.bias",
             "fc3.weight", "fc3.bias", "fc4.weight", "fc4.bias")


def _ptr(t):
    if t is None:
        return None
    return C.c_void_p(t.data_ptr())


def nccl_comm_ptr(group, device):
    """ncclComm_t of ``group`` as an integer (torch owns it; never free it)."""
    backend = group._get_backend(torch.device("cuda", device.index))
    if not hasattr(backend, "_comm_ptr"):
        raise RuntimeError("this torch build does not expose ProcessGroupNCCL._comm_ptr()")
    try:
        return int(backend._comm_ptr())
    except Exception:
        # communicator not created yet (process group made without device_id): one tiny collective creates it
        t = torch.zeros(1, device=device)
        dist.all_reduce(t, group=group)
        torch.cuda.synchronize(device)
        return int(backend._comm_ptr())


class VAEEngine:
    def __init__(self, in_dim=784, hidden=400, latent=20, max_batch=128, gemm_mode="tf32x3", use_graph=True,
                 group=None, device=None, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, exchange=None):
        """``exchange``: how a group of n > 1 ranks sums its gradients -- "auto" (peer memory for groups of up to
        MDT_P2P_MAX_RANKS=4 ranks on one node, else NCCL), "nccl", "p2p" (one-shot) or "p2p_two_shot"."""
        if not torch.cuda.is_available():
            raise RuntimeError("multidisttorch_b200.VAEEngine needs a CUDA device; there is no CPU path")
        self.lib = _lib.load()
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.D, self.H, self.Z, self.max_batch = in_dim, hidden, latent, max_batch
        self.gemm_mode = gemm_mode
        self.group = group
        self.group_size = dist.get_world_size(group) if (group is not None and dist.is_initialized()) else 1
        self.group_rank = dist.get_rank(group) if self.group_size > 1 else 0
        if exchange is None:      # environment default (MDT_P2P_ALLREDUCE=0 keeps NCCL everywhere)
            exchange = {"0": "nccl"}.get(os.environ.get("MDT_P2P_ALLREDUCE", "1"), "auto")
        self.exchange = exchange
        self.cfg = mdt_cfg(in_dim, hidden, latent, max_batch, MDT_GEMM[gemm_mode], int(bool(use_graph)),
                           MDT_EXCHANGE[exchange], 0)
        off = (C.c_int64 * 8)()
        num = (C.c_int64 * 8)()
        tot = C.c_int64()
        check(self.lib.mdt_param_layout(C.byref(self.cfg), off, num, C.byref(tot)))
        self.offsets = dict(zip(SEGMENTS, (int(o) for o in off)))
        self.numels = dict(zip(SEGMENTS, (int(n) for n in num)))
        self.total = int(tot.value)
        self.shapes = {"W1": (hidden, in_dim), "b1": (hidden,), "W2x": (2 * latent, hidden), "b2x": (2 * latent,),
                       "W3": (hidden, latent), "b3": (hidden,), "W4": (in_dim, hidden), "b4": (in_dim,)}
        self._h = None
        with torch.cuda.device(self.device):
            z = lambda: torch.zeros(self.total, dtype=torch.float32, device=self.device)
            self.params, self.grads, self.exp_avg, self.exp_avg_sq = z(), z(), z(), z()
            self._loss = torch.zeros(1, dtype=torch.float32, device=self.device)
            comm = None
            if self.group_size > 1:
                comm = C.c_void_p(nccl_comm_ptr(group, self.device))
            h = C.c_void_p()
            torch.cuda.synchronize(self.device)
            check(self.lib.mdt_engine_create(C.byref(self.cfg), comm, self.group_size, self.group_rank, None, C.byref(h)))
            self._h = h
            check(self.lib.mdt_engine_bind_buffers(h, _ptr(self.params), _ptr(self.grads), _ptr(self.exp_avg),
                                                   _ptr(self.exp_avg_sq), self.total))
            s = C.c_void_p()
            check(self.lib.mdt_engine_stream(h, C.byref(s)))
            self.stream = torch.cuda.ExternalStream(s.value, device=self.device)
            self.p2p = False
            # Gradient exchange over NVLink peer memory inside the Adam kernel (one-shot, or two-shot when a rank would
            # read more than 16 MB from its peers) against NCCL: measured figures in DESIGN.md section 5.
            max_ranks = int(os.environ.get("MDT_P2P_MAX_RANKS", "4"))
            want = self.group_size <= max_ranks or exchange in ("p2p", "p2p_two_shot")
            if self.group_size > 1 and gemm_mode != "fp32" and exchange != "nccl" and want and self.group_size <= 8:
                self.p2p = self._connect_peers()
        self.set_adam(lr, betas, eps, 0)

    def set_sm_share(self, n_sharing):
        """Size this engine's kernels for 1/n_sharing of the SMs (several trials on one GPU; mdt_engine_set_sm_share)."""
        check(self.lib.mdt_engine_set_sm_share(self._h, int(n_sharing)))

    def exchange_status(self):
        """(enabled, timed_out) of the peer-memory gradient exchange (mdt_engine_exchange_status)."""
        en, to = C.c_int(), C.c_int()
        check(self.lib.mdt_engine_exchange_status(self._h, C.byref(en), C.byref(to)))
        return bool(en.value), bool(to.value)

    def _connect_peers(self):
        """Exchange the CUDA IPC handles of the group's gradient-exchange regions (include/mdt_b200.h,
        mdt_engine_set_peers).  Every rank of the group must agree, so the outcome is all-reduced: if any rank
        cannot map its peers (ranks on different nodes, no peer access) the whole group stays on NCCL."""
        mine = (C.c_ubyte * 64)()
        ok = self.lib.mdt_engine_exchange_handle(self._h, mine) == 0
        handles = [None] * self.group_size
        dist.all_gather_object(handles, bytes(mine) if ok else None, group=self.group)
        ok = ok and all(h is not None for h in handles)
        if ok:
            blob = (C.c_ubyte * (64 * self.group_size)).from_buffer_copy(b"".join(handles))
            ok = self.lib.mdt_engine_set_peers(self._h, blob, self.group_size) == 0
        flag = torch.tensor([1 if ok else 0], device=self.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.group)
        if int(flag.item()) == 0:
            if ok:   # this rank could map its peers, another rank of the group could not: everybody stays on NCCL
                check(self.lib.mdt_engine_disable_peers(self._h))
            return False
        return True

    # ------------------------------------------------------------------ buffers
    def segment(self, buf, name):
        o, n = self.offsets[name], self.numels[name]
        return buf[o:o + n].view(self.shapes[name])

    def named_parameter_views(self, buf=None):
        """Reference-named views (fc1.weight ... fc4.bias) into a flat buffer."""
        buf = self.params if buf is None else buf
        Z = self.Z
        seg = lambda n: self.segment(buf, n)
        return {
            "fc1.weight": seg("W1"), "fc1.bias": seg("b1"),
            "fc21.weight": seg("W2x")[:Z], "fc21.bias": seg("b2x")[:Z],
            "fc22.weight": seg("W2x")[Z:], "fc22.bias": seg("b2x")[Z:],
            "fc3.weight": seg("W3"), "fc3.bias": seg("b3"),
            "fc4.weight": seg("W4"), "fc4.bias": seg("b4"),
        }

    def load_segments(self, seg_dict):
        """seg_dict: segment name -> array/tensor (flat-layout naming)."""
        with torch.no_grad():
            for k in SEGMENTS:
                v = seg_dict[k]
                src = v if torch.is_tensor(v) else torch.from_numpy(np.ascontiguousarray(v))
                self.segment(self.params, k).copy_(src.to(self.device, torch.float32))
        self.params_changed()

    def load_reference_state_dict(self, sd):
        with torch.no_grad():
            views = self.named_parameter_views()
            for k in REF_NAMES:
                views[k].copy_(sd[k].to(self.device, torch.float32))
        self.params_changed()

    def segments_to_numpy(self, buf=None):
        buf = self.params if buf is None else buf
        self.synchronize()
        torch.cuda.synchronize(self.device)
        return {k: self.segment(buf, k).detach().cpu().numpy().copy() for k in SEGMENTS}

    def params_changed(self):
        """Call after writing ``params`` from outside the engine: re-binds so derived
        copies (transposed weight shadows of the tcgen05 path) are rebuilt."""
        torch.cuda.synchronize(self.device)
        check(self.lib.mdt_engine_bind_buffers(self._h, _ptr(self.params), _ptr(self.grads), _ptr(self.exp_avg),
                                               _ptr(self.exp_avg_sq), self.total))

    def broadcast_params(self):
        check(self.lib.mdt_engine_broadcast_params(self._h))

    # ------------------------------------------------------------------ state
    def set_adam(self, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, completed_steps=0):
        h = mdt_adam(lr, betas[0], betas[1], eps)
        check(self.lib.mdt_engine_set_adam(self._h, C.byref(h), completed_steps))

    def set_sgd(self, lr, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False, completed_steps=0):
        """optim.SGD in place of Adam (mdt_engine_set_sgd); exp_avg becomes the momentum buffer."""
        h = mdt_sgd(lr, momentum, dampening, weight_decay, int(bool(nesterov)), 0)
        check(self.lib.mdt_engine_set_sgd(self._h, C.byref(h), completed_steps))

    def set_rng(self, seed, offset=0):
        check(self.lib.mdt_engine_set_rng(self._h, seed, offset))

    def state(self):
        t, o = C.c_int64(), C.c_uint64()
        check(self.lib.mdt_engine_get_state(self._h, C.byref(t), C.byref(o)))
        return int(t.value), int(o.value)

    def synchronize(self):
        self.stream.synchronize()

    def input_buffer_ptr(self):
        p = C.c_void_p()
        check(self.lib.mdt_engine_input_buffer(self._h, C.byref(p)))
        return p.value

    def workspace(self, name, rows):
        """Copy of an internal activation buffer (tests/debugging)."""
        p, ld = C.c_void_p(), C.c_int()
        check(self.lib.mdt_engine_workspace(self._h, name.encode(), C.byref(p), C.byref(ld)))
        self.synchronize()
        view = _DevView(p.value, (rows, ld.value), self.device)
        return torch.as_tensor(view, device=self.device).clone()

    # ------------------------------------------------------------------ the step
    def _x2d(self, x):
        x = x.reshape(-1, self.D)
        assert x.dtype == torch.float32 and x.is_contiguous()
        return x

    def train_step(self, x, eps=None, want_loss=True):
        """One ref:vae-hpo.py:68-74 step on device-resident x.  Returns the loss as a
        1-element device tensor (no host sync) or None."""
        x = self._x2d(x)
        check(self.lib.mdt_train_step(self._h, _ptr(x), x.shape[0], _ptr(eps), _ptr(self._loss) if want_loss else None))
        return self._loss if want_loss else None

    def train_steps(self, pool, first_batch, n_steps, losses=None):
        """``n_steps`` consecutive steps on a device-resident pool [n_batches, B, D] (mdt_train_steps): step i trains on
        ``pool[(first_batch + i) % n_batches]``.  ``losses``: optional device tensor of n_steps floats."""
        nb, B = pool.shape[0], pool.shape[1]
        assert pool.dtype == torch.float32 and pool.is_contiguous() and pool.numel() == nb * B * self.D
        src = mdt_batch_source(pool.data_ptr(), None, None, nb, 0)
        check(self.lib.mdt_train_steps(self._h, C.byref(src), B, first_batch, n_steps, _ptr(losses)))

    def train_steps_u8(self, images_u8, indices, batch, first_batch, n_steps, losses=None):
        """Same from a device-resident uint8 image table [n_images, D] and the sampler's int64 index order
        (batch-major, n_batches * batch entries): gather + /255 run inside each step's first kernel."""
        assert images_u8.dtype == torch.uint8 and images_u8.is_contiguous() and indices.dtype == torch.int64
        nb = indices.numel() // batch
        src = mdt_batch_source(None, images_u8.data_ptr(), indices.data_ptr(), nb, images_u8.shape[0])
        check(self.lib.mdt_train_steps(self._h, C.byref(src), batch, first_batch, n_steps, _ptr(losses)))

    def exchange_bench(self, iters=20):
        """Seconds per gradient exchange + optimizer pass of this group (mdt_engine_exchange_bench; scratch engines only)."""
        s = C.c_double()
        check(self.lib.mdt_engine_exchange_bench(self._h, iters, C.byref(s)))
        return float(s.value)

    def train_step_staged(self, batch, eps=None, want_loss=True):
        """Step on a batch that is already in the engine's stage buffer (device-resident data path: no copy)."""
        check(self.lib.mdt_train_step(self._h, C.c_void_p(self.input_buffer_ptr()), batch, _ptr(eps),
                                      _ptr(self._loss) if want_loss else None))
        return self._loss if want_loss else None

    def train_step_host(self, x_host, eps_host=None):
        """Same step with HOST tensors (pinned for async copies); returns float loss."""
        x = self._x2d(x_host)
        out = C.c_float()
        check(self.lib.mdt_train_step_host(self._h, _ptr(x), x.shape[0], _ptr(eps_host), C.byref(out)))
        return float(out.value)

    def submit_host(self, x_host, eps_host=None):
        """First half of train_step_host: enqueue the step, return a ticket.  ``x_host`` must stay untouched until
        ``wait_host(ticket)`` has returned (pinned memory is read in place by the device)."""
        x = self._x2d(x_host)
        t = C.c_uint32()
        check(self.lib.mdt_train_step_host_submit(self._h, _ptr(x), x.shape[0], _ptr(eps_host), C.byref(t)))
        self._inflight = getattr(self, "_inflight", {})
        self._inflight[t.value] = (x, eps_host)          # keep the buffers alive while the device may read them
        return t.value

    def submit_host_chunk(self, x_host):
        """mdt_train_steps_host_submit: ``x_host`` = [n, B, D] host batches (n <= 16, pinned for an asynchronous copy);
        returns the tickets of the n steps.  Keep ``x_host`` untouched until the first ticket has been waited for."""
        n, B = x_host.shape[0], x_host.shape[1]
        assert x_host.dtype == torch.float32 and x_host.is_contiguous() and x_host.numel() == n * B * self.D
        t = C.c_uint32()
        check(self.lib.mdt_train_steps_host_submit(self._h, _ptr(x_host), B, n, C.byref(t)))
        self._inflight = getattr(self, "_inflight", {})
        self._inflight[t.value] = (x_host, None)
        return [(t.value + i) & 0x7fffffff for i in range(n)]

    def wait_host(self, ticket):
        """Second half: the loss of the step ``submit_host`` returned ``ticket`` for."""
        out = C.c_float()
        check(self.lib.mdt_train_step_host_wait(self._h, C.c_uint32(ticket), C.byref(out)))
        getattr(self, "_inflight", {}).pop(ticket, None)
        return float(out.value)

    def forward(self, x, eps=None):
        x = self._x2d(x)
        B = x.shape[0]
        recon = torch.empty(B, self.D, dtype=torch.float32, device=self.device)
        mu = torch.empty(B, self.Z, dtype=torch.float32, device=self.device)
        lv = torch.empty(B, self.Z, dtype=torch.float32, device=self.device)
        check(self.lib.mdt_forward(self._h, _ptr(x), B, _ptr(eps), _ptr(recon), _ptr(mu), _ptr(lv)))
        return recon, mu, lv

    def backward(self, d_recon, d_mu, d_logvar):
        check(self.lib.mdt_backward(self._h, _ptr(d_recon.contiguous()), _ptr(d_mu.contiguous()),
                                    _ptr(d_logvar.contiguous())))

    def allreduce_grads(self):
        check(self.lib.mdt_allreduce_grads(self._h))

    def adam_step(self, grad_scale=None):
        gs = 1.0 / self.group_size if grad_scale is None else grad_scale
        check(self.lib.mdt_adam_step(self._h, gs))

    def eval_step(self, x, eps=None, want_recon=False):
        x = self._x2d(x)
        recon = torch.empty(x.shape[0], self.D, dtype=torch.float32, device=self.device) if want_recon else None
        check(self.lib.mdt_eval_step(self._h, _ptr(x), x.shape[0], _ptr(eps), _ptr(self._loss), _ptr(recon)))
        return self._loss, recon

    def decode(self, z):
        z = z.reshape(-1, self.Z).contiguous()
        recon = torch.empty(z.shape[0], self.D, dtype=torch.float32, device=self.device)
        check(self.lib.mdt_decode(self._h, _ptr(z), z.shape[0], _ptr(recon)))
        return recon

    def launch_count(self):
        n = C.c_int64()
        check(self.lib.mdt_engine_launch_count(self._h, C.byref(n)))
        return int(n.value)

    def profile_step(self, x, iters=20):
        """Per-kernel device time of the (un-graphed) step: list of (label, avg_ms)."""
        x = self._x2d(x)
        buf = C.create_string_buffer(8192)
        check(self.lib.mdt_engine_profile_step(self._h, _ptr(x), x.shape[0], iters, buf, len(buf)))
        rows = [l.split() for l in buf.value.decode().splitlines() if l.strip()]
        return [(n, float(ms)) for n, ms in rows]

    def close(self):
        if getattr(self, "_h", None) is not None:
            self.lib.mdt_engine_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class _DevView:
    """Expose a raw device pointer to torch through __cuda_array_interface__."""

    def __init__(self, ptr, shape, device):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f4", "data": (int(ptr), False),
                                         "version": 2, "strides": None}
        self.device = device


# ---------------------------------------------------------------------- stand-alone ops
def _cur_stream(dev):
    return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def train_step_many(engines, xs):
    """One step of several independent trials on one GPU in a single C-ABI call; xs[i] is the
    [B, D] device batch of engines[i]."""
    n = len(engines)
    hs = (C.c_void_p * n)(*[e._h for e in engines])
    ps = (C.c_void_p * n)(*[x.data_ptr() for x in xs])
    check(_lib.load().mdt_train_step_many(hs, ps, n, xs[0].shape[0]))


class TrialGroup:
    """T single-rank trials of one shape on one GPU, stepped in lock-step with ONE launch per layer (mdt_group_*,
    SURVEY 8f-3).  The engines stay usable on their own; each keeps its parameters, state and losses."""

    def __init__(self, engines):
        self.engines = list(engines)
        self.lib = _lib.load()
        n = len(self.engines)
        hs = (C.c_void_p * n)(*[e._h for e in self.engines])
        g = C.c_void_p()
        check(self.lib.mdt_group_create(hs, n, C.byref(g)))
        self._g = g

    def train_steps(self, pools, first_batch, n_steps):
        """pools: one [n_batches, B, D] device tensor per trial (the same tensor may be given to several trials)."""
        n = len(self.engines)
        assert len(pools) == n
        B = pools[0].shape[1]
        srcs = (mdt_batch_source * n)()
        for t, pl in enumerate(pools):
            assert pl.dtype == torch.float32 and pl.is_contiguous() and pl.shape[1] == B
            srcs[t] = mdt_batch_source(pl.data_ptr(), None, None, pl.shape[0], 0)
        self._keep = pools
        check(self.lib.mdt_group_train_steps(self._g, srcs, B, first_batch, n_steps))

    def synchronize(self):
        self.engines[0].synchronize()

    def close(self):
        if getattr(self, "_g", None) is not None:
            self.lib.mdt_group_destroy(self._g)
            self._g = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def adam_flat(p, g, m, v, step, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, grad_scale=1.0):
    h = mdt_adam(lr, betas[0], betas[1], eps)
    check(_lib.load().mdt_adam_flat(_ptr(p), _ptr(g), _ptr(m), _ptr(v), p.numel(), C.byref(h), step, grad_scale,
                                    _cur_stream(p.device)))


def reparam(mu, logvar, eps=None, seed=0, offset=0):
    B, Z = mu.shape
    z = torch.empty(B, Z, dtype=torch.float32, device=mu.device)
    eps_out = torch.empty(B, Z, dtype=torch.float32, device=mu.device)
    assert mu.stride(0) == logvar.stride(0) and mu.stride(1) == 1 and logvar.stride(1) == 1
    check(_lib.load().mdt_reparam(_ptr(mu), _ptr(logvar), mu.stride(0), _ptr(eps), seed, offset, B, Z, _ptr(z), Z,
                                  _ptr(eps_out), _cur_stream(mu.device)))
    return z, eps_out


def loss_bce_kl(recon, x, mu, logvar, want_grads=True):
    B, D = recon.shape
    Z = mu.shape[1]
    loss = torch.empty(1, dtype=torch.float32, device=recon.device)
    dr = torch.empty_like(recon) if want_grads else None
    dm = torch.empty_like(mu) if want_grads else None
    dl = torch.empty_like(logvar) if want_grads else None
    check(_lib.load().mdt_loss_bce_kl(_ptr(recon), _ptr(x), _ptr(mu), _ptr(logvar), B, D, Z, _ptr(loss), _ptr(dr),
                                      _ptr(dm), _ptr(dl), _cur_stream(recon.device)))
    return loss, dr, dm, dl


def loss_bce_kl_noalloc(recon, x, mu, logvar, loss, d_recon=None, d_mu=None, d_logvar=None):
    """mdt_loss_bce_kl_noalloc: every buffer, outputs included, is supplied by the caller, and the
    library keeps its scratch in a fixed in-image array, so nothing is allocated anywhere on the
    way (not by torch either).  For callers that hold a fully committed memory budget or cannot
    afford cudaMalloc's device-wide synchronisation.  batch*in_dim <= 1,048,576."""
    B, D = recon.shape
    check(_lib.load().mdt_loss_bce_kl_noalloc(_ptr(recon), _ptr(x), _ptr(mu), _ptr(logvar), B, D, mu.shape[1],
                                              _ptr(loss), _ptr(d_recon), _ptr(d_mu), _ptr(d_logvar),
                                              _cur_stream(recon.device)))
    return loss


def linear(x, w, b=None, act="none", gemm_mode="fp32"):
    M, K = x.shape
    N = w.shape[0]
    y = torch.empty(M, N, dtype=torch.float32, device=x.device)
    check(_lib.load().mdt_linear(_ptr(x), _ptr(w), _ptr(b), M, N, K, _lib.MDT_ACT[act], MDT_GEMM[gemm_mode], _ptr(y),
                                 _cur_stream(x.device)))
    return y
