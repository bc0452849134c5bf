"""Per-trial-group gradient averaging on the host API level.

DDP's default hook divides by the group size and then SUM-all-reduces
(ref:vae-hpo.py:130).  The B200 engine keeps all gradients in
one flat buffer, so one group = one (or two, bucketed) collectives per step,
and the 1/n factor is not applied here at all on the CUDA path: it is folded
into the fused Adam kernel (``grad_scale``).  ``average_flat_`` is the
torch.distributed leg used for groups whose communicator lives in torch
(gloo in the CPU plumbing tests; NCCL when the C-ABI is not handed a comm)."""
import torch
import torch.distributed as dist


def allreduce_sum_flat_(flat: torch.Tensor, group) -> torch.Tensor:
    """SUM over the ranks of ``group`` in place; no scaling (the Adam kernel
    multiplies by 1/n)."""
    if group is not None and dist.get_world_size(group) > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    return flat


def average_flat_(flat: torch.Tensor, group) -> torch.Tensor:
    """DDP semantics on a flat buffer: g <- sum_r (g_r / n)."""
    n = dist.get_world_size(group) if group is not None else 1
    if n > 1:
        flat.div_(n)
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    return flat
