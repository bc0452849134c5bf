"""HPO hyper-parameter surface of a trial group.

The reference varies only the number of epochs between its concurrent trials (``run(gid, batch_size, gid*2+1, group)``,
ref:vae-hpo.py:202) and every rank writes into ``results-{rank}`` (ref:vae-hpo.py:107-118, 166-170), so the ranks of
one group collide with nothing but also share nothing.  Here every trial group can get its own optimizer
hyper-parameters and batch size from the environment, without touching the scripts:

    MDT_HPO_LR=1e-3,3e-4  MDT_HPO_BETA1=0.9,0.8  MDT_HPO_BETA2=0.999  MDT_HPO_EPS=1e-8  MDT_HPO_BATCH=128,256

Each variable is a comma-separated list indexed by the trial's group id (wrapping around), so one value applies to all
trials.  ``multidisttorch_b200.dropin`` consults :func:`trial_config` when it stands in for ``optim.Adam`` of a trial
(lr, betas, eps) and for the trial's training ``DataLoader`` (batch size); callers of the engine API use it directly.
(Hidden width and epoch count are fixed by the script's own ``class VAE`` and ``for epoch`` loop, ref:vae-hpo.py:20-27,
159: nothing outside the script can vary them, so they are not part of this surface.)
"""
import os
from dataclasses import dataclass

import torch.distributed as dist

_GROUP_INDEX = {}   # id(ProcessGroup) -> index in the list setup_ddp_groups returned


def register_groups(groups):
    """Called by ``utils.setup_ddp_groups``: remember which list position (= trial / group id) each group has."""
    for gid, g in enumerate(groups):
        _GROUP_INDEX[id(g)] = gid


def group_index(group):
    """Trial (group) id of ``group`` as created by ``setup_ddp_groups``; None for the world group or unknown groups."""
    if group is None:
        return None
    return _GROUP_INDEX.get(id(group))


def my_group_index(groups):
    """Position of the group this rank belongs to (the reference's ``group_id`` loop, ref:vae-hpo.py:196-202)."""
    for gid, g in enumerate(groups):
        if dist.get_rank(g) >= 0:
            return gid
    return None


@dataclass(frozen=True)
class TrialConfig:
    group_id: int
    lr: float = 1e-3
    beta1: float = 0.9
    beta2: float = 0.999
    eps: float = 1e-8
    batch_size: int = 128

    @property
    def betas(self):
        return (self.beta1, self.beta2)


_FIELDS = {"lr": ("MDT_HPO_LR", float), "beta1": ("MDT_HPO_BETA1", float), "beta2": ("MDT_HPO_BETA2", float),
           "eps": ("MDT_HPO_EPS", float), "batch_size": ("MDT_HPO_BATCH", int)}


def _pick(raw, conv, gid, name):
    vals = [v.strip() for v in raw.split(",") if v.strip()]
    if not vals:
        raise ValueError(f"{name} is set but empty")
    try:
        return conv(vals[gid % len(vals)])
    except ValueError as e:
        raise ValueError(f"{name}={raw!r}: {e}") from None


def trial_config(group_id, env=None, **defaults):
    """Hyper-parameters of trial ``group_id``: ``defaults`` (what the script asked for) overridden by MDT_HPO_*."""
    env = os.environ if env is None else env
    gid = int(group_id or 0)
    kw = dict(defaults)
    for field, (name, conv) in _FIELDS.items():
        raw = env.get(name)
        if raw is not None:
            kw[field] = _pick(raw, conv, gid, name)
    cfg = TrialConfig(group_id=gid, **kw)
    if not (cfg.lr > 0 and 0 <= cfg.beta1 < 1 and 0 <= cfg.beta2 < 1 and cfg.eps > 0):
        raise ValueError(f"invalid Adam hyper-parameters for trial {gid}: {cfg}")
    if cfg.batch_size <= 0:
        raise ValueError(f"invalid batch size for trial {gid}: {cfg}")
    return cfg


def overridden(env=None):
    """True if any MDT_HPO_* variable is set."""
    env = os.environ if env is None else env
    return any(name in env for name, _ in _FIELDS.values())
