"""Rank-to-group mapping and subgroup bootstrap -- the kept host API.

Same names, argument meaning, return values, prints and error behaviour as the
reference's ``utils.py`` (ref:utils.py:1-174), so ``example-subgroup.py`` and
``vae-hpo.py`` run unchanged on top of it (``from utils import *`` resolves to
the shim at the repository root, which re-exports this module).

What is different underneath, B200-first:

* ``setup_ddp`` binds one GPU per rank *before* the process group exists and
  passes ``device_id`` to ``init_process_group`` so the world NCCL communicator
  is created eagerly and every ``dist.new_group`` becomes an ``ncclCommSplit``
  over NVLink/NVSwitch (the reference leaves the device unbound, which makes
  each subgroup a lazily bootstrapped ``ncclCommInitRank``).
* on a CUDA box the world group carries a gloo leg for CPU tensors
  (``cpu:gloo,cuda:nccl``) because ref:example-subgroup.py:25-33 all-gathers CPU
  int64 tensors.
* importing the module installs the drop-in hooks of
  ``multidisttorch_b200.dropin`` (CUDA only, ``MDT_DROPIN=0`` disables).

The rank -> group assignment (``setup_ddp_groups``) is integer-exact with the
reference; tests/test_utils_api.py pins it against tests/golden/api.json, which
the reference itself produced.
"""
import os
import re
import socket

import psutil
import torch
import torch.distributed as dist

_LAUNCHER_ENV = (
    ("OMPI_COMM_WORLD_SIZE", "OMPI_COMM_WORLD_RANK"),   # mpirun / jsrun (Summit), ref:utils.py:13-16
    ("SLURM_NPROCS", "SLURM_PROCID"),                   # srun (CADES), ref:utils.py:17-20
)


def init_comm_size_and_rank():
    """(world_size, world_rank) from the launcher's environment; (1, 0) when no
    launcher is detected.  torchrun's RANK/WORLD_SIZE are deliberately NOT
    consulted, like ref:utils.py:9-26."""
    for size_key, rank_key in _LAUNCHER_ENV:
        size, rank = os.getenv(size_key), os.getenv(rank_key)
        if size and rank:
            return int(size), int(rank)
    return 1, 0


def get_comm_size_and_rank():
    """(world_size, world_rank) of the initialised default group, else (1, 0)
    (ref:utils.py:28-38)."""
    if not dist.is_initialized():
        return 1, 0
    return int(dist.get_world_size()), int(dist.get_rank())


def find_ifname(myaddr):
    """Name of the network interface that owns ``myaddr`` (e.g. "lo" for
    127.0.0.1), or None.  Used for GLOO_SOCKET_IFNAME (ref:utils.py:40-56)."""
    wanted = socket.gethostbyname(myaddr)
    for nic, addrs in psutil.net_if_addrs().items():
        if any(a.address == wanted for a in addrs):
            return nic
    return None


_NODE_BLOCK = re.compile(r"([\w-]+)((?:\[[\d\-,]+\])*)")


def parse_slurm_nodelist(nodelist):
    """Expand a SLURM_NODELIST string, keeping zero padding
    (ref:utils.py:59-90):  "g[05,07-08,13],h[01,12]" ->
    [g05, g07, g08, g13, h01, h12]."""
    nodes = []
    for prefix, brackets in _NODE_BLOCK.findall(nodelist):
        if not brackets:
            nodes.append(prefix)
            continue
        # the reference's greedy "\[(.*)\]" spans from the first "[" to the last "]"
        inner = brackets[1:-1]
        for item in inner.split(","):
            if "-" not in item:
                nodes.append(prefix + item)
                continue
            # a malformed range raises AttributeError here, as it does in the reference
            lo, hi = re.match(r"(\d+)-(\d+)", item).groups()
            width = len(lo)
            nodes.extend(prefix + str(i).zfill(width) for i in range(int(lo), int(hi) + 1))
    return nodes


def _pick_backend():
    forced = os.getenv("DDP_BACKEND")
    if forced is not None:
        return forced
    if dist.is_nccl_available() and torch.cuda.is_available():
        return "nccl"
    if dist.is_gloo_available():
        return "gloo"
    raise RuntimeError("No parallel backends available")


def _pick_master():
    addr = os.getenv("MASTER_ADDR", "127.0.0.1")
    port = os.getenv("MASTER_PORT", "8889")
    if os.getenv("LSB_HOSTS") is not None:            # Summit LSF
        addr = os.environ["LSB_HOSTS"].split()[1]
    elif os.getenv("LSB_MCPU_HOSTS") is not None:
        addr = os.environ["LSB_MCPU_HOSTS"].split()[2]
    elif os.getenv("SLURM_NODELIST") is not None:     # CADES
        addr = parse_slurm_nodelist(os.environ["SLURM_NODELIST"])[0]
    return addr, port


def local_device_index(world_rank):
    """GPU this rank drives: the launcher's local rank when it exports one,
    else round-robin over the visible devices."""
    for key in ("LOCAL_RANK", "OMPI_COMM_WORLD_LOCAL_RANK", "SLURM_LOCALID", "MV2_COMM_WORLD_LOCAL_RANK"):
        val = os.getenv(key)
        if val:
            return int(val) % max(torch.cuda.device_count(), 1)
    return world_rank % max(torch.cuda.device_count(), 1)


def setup_ddp():
    """Initialise the world process group from launcher env vars and return
    (world_size, world_rank) -- ref:utils.py:93-144, same prints."""
    backend = _pick_backend()
    world_size, world_rank = init_comm_size_and_rank()
    master_addr, master_port = _pick_master()

    try:
        if backend in ("nccl", "gloo"):
            os.environ.update(MASTER_ADDR=master_addr, MASTER_PORT=master_port,
                              WORLD_SIZE=str(world_size), RANK=str(world_rank))
        if backend == "gloo" and "GLOO_SOCKET_IFNAME" not in os.environ:
            ifname = find_ifname(master_addr)
            if ifname is not None:
                os.environ["GLOO_SOCKET_IFNAME"] = ifname

        print("Distributed data parallel: %s master at %s:%s" % (backend, master_addr, master_port))

        if not dist.is_initialized():
            if backend == "nccl" and torch.cuda.is_available():
                dev = torch.device("cuda", local_device_index(world_rank))
                torch.cuda.set_device(dev)
                if "GLOO_SOCKET_IFNAME" not in os.environ:
                    ifname = find_ifname(master_addr)
                    if ifname is not None:
                        os.environ["GLOO_SOCKET_IFNAME"] = ifname
                dist.init_process_group(backend="cpu:gloo,cuda:nccl", init_method="env://", device_id=dev)
            else:
                dist.init_process_group(backend=backend, init_method="env://")
    except KeyError:
        print("DDP has to be initialized within a job - Running in sequential mode")

    return world_size, world_rank


def group_ranks(world_size, num_groups):
    """The rank lists of ``setup_ddp_groups``: group g owns the contiguous
    ranks [k*g, k*g+k), k = world_size // num_groups; ranks >= k*num_groups
    belong to no group (ref:utils.py:152-156)."""
    k = world_size // num_groups
    return [list(range(k * g, k * g + k)) for g in range(num_groups)]


def setup_ddp_groups(num_groups):
    """Create ``num_groups`` disjoint subgroups; EVERY rank creates EVERY
    group, in order, and gets the full list back (ref:utils.py:146-163)."""
    world_size, world_rank = get_comm_size_and_rank()
    print("world_size, world_rank:", world_size, world_rank)
    assert world_size >= num_groups, (
        f"Number of groups {num_groups} requested exceeds number of total processes {world_size} available")

    groups = [dist.new_group(ranks=ranks) for ranks in group_ranks(world_size, num_groups)]
    from . import hpo
    hpo.register_groups(groups)        # trial id of each group, for the per-trial hyper-parameter surface (hpo.py)
    for gid, grp in enumerate(groups):
        if dist.get_rank(grp) >= 0:
            print(f"Rank {world_rank} is in group {gid}")
    return groups


def print0(*args, sep=" ", process_group=None):
    """Print "[world_rank:group_rank] msg" on group-rank 0 of ``process_group``
    only (ref:utils.py:165-174)."""
    grank = dist.get_rank(process_group)
    if grank == 0:
        print(f"[{dist.get_rank()}:{grank}]", sep.join(str(a) for a in args))


def _install_dropin():
    if os.getenv("MDT_DROPIN", "1") == "0" or not torch.cuda.is_available():
        return
    from . import dropin
    dropin.install()


_install_dropin()
