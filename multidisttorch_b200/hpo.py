"""HPO hyper-parameter surface of a trial group.

The reference varies only the number of epochs between its concurrent trials (``run(gid, batch_size, gid*2+1, group)``,
ref:vae-hpo.py:202) and every rank writes into ``results-{rank}`` (ref:vae-hpo.py:107-118, 166-170), so the ranks of
one group collide with nothing but also share nothing.  Here every trial group can get its own optimizer
hyper-parameters, batch size, hidden width and epoch count from the environment, without touching the scripts:

    MDT_HPO_LR=1e-3,3e-4  MDT_HPO_BETA1=0.9,0.8  MDT_HPO_BETA2=0.999  MDT_HPO_EPS=1e-8
    MDT_HPO_BATCH=128,256  MDT_HPO_HIDDEN=400  MDT_HPO_EPOCHS=1,3

Each variable is a comma-separated list indexed by the trial's group id (wrapping around), so one value applies to all
trials.  ``multidisttorch_b200.dropin`` consults :func:`trial_config` when it stands in for ``optim.Adam`` of a trial;
callers of the engine API use it directly.  :func:`results_dir` gives one directory per (group, rank-in-group).
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
    hidden: int = 400
    epochs: int = 1

    @property
    def betas(self):
        return (self.beta1, self.beta2)


_FIELDS = {"lr": ("MDT_HPO_LR", float), "beta1": ("MDT_HPO_BETA1", float), "beta2": ("MDT_HPO_BETA2", float),
           "eps": ("MDT_HPO_EPS", float), "batch_size": ("MDT_HPO_BATCH", int), "hidden": ("MDT_HPO_HIDDEN", int),
           "epochs": ("MDT_HPO_EPOCHS", int)}


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
    if cfg.batch_size <= 0 or cfg.hidden <= 0 or cfg.hidden % 4 or cfg.epochs < 0:
        raise ValueError(f"invalid shape hyper-parameters for trial {gid}: {cfg} (hidden must be a multiple of 4)")
    return cfg


def overridden(env=None):
    """True if any MDT_HPO_* variable is set."""
    env = os.environ if env is None else env
    return any(name in env for name, _ in _FIELDS.values())


def results_dir(group_id, group_rank=0, root=".", create=True):
    """``results-g{group}`` for group-rank 0, ``results-g{group}-r{rank}`` for the others: one directory per trial
    member instead of the reference's world-rank keyed ``results-{rank}``."""
    name = f"results-g{int(group_id)}" + ("" if int(group_rank) == 0 else f"-r{int(group_rank)}")
    path = os.path.join(root, name)
    if create:
        os.makedirs(path, exist_ok=True)
    return path
