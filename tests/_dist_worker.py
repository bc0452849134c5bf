"""One rank of a CPU/gloo multi-process test.  Launched by tests/dist_launch.py
with OMPI_COMM_WORLD_SIZE/RANK set, like mpirun would (ref:utils.py:13-16)."""
import io
import json
import os
import sys
from contextlib import redirect_stdout

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from utils import *  # noqa: E402,F401,F403  (the repo-root drop-in shim)


def main():
    mode, ngroups, out_dir = sys.argv[1], int(sys.argv[2]), sys.argv[3]
    buf = io.StringIO()
    with redirect_stdout(buf):
        size, rank = setup_ddp()
        groups = setup_ddp_groups(ngroups) if ngroups > 0 else []
    res = {"size": size, "rank": rank, "stdout": buf.getvalue(),
           "membership": [int(dist.get_rank(g)) for g in groups],
           "get_comm": list(get_comm_size_and_rank())}
    from multidisttorch_b200 import hpo
    res["hpo_group_index"] = [hpo.group_index(g) for g in groups]
    res["hpo_my_group"] = hpo.my_group_index(groups)
    mine = [g for g in groups if dist.get_rank(g) >= 0]
    if mode == "gather" and mine:
        # what ref:example-subgroup.py:25-33 does, on the group this rank belongs to
        grp = mine[0]
        n = dist.get_world_size(grp)
        got = [torch.tensor([0]) for _ in range(n)]
        dist.all_gather(got, torch.tensor([rank]), group=grp)
        res["gather"] = [int(t.item()) for t in got]
        pb = io.StringIO()
        with redirect_stdout(pb):
            print0("hello", 3, process_group=grp)
        res["print0"] = pb.getvalue()
    if mode == "allreduce" and mine:
        # per-group gradient averaging with DDP's semantics on a flat buffer: g <- sum_r (g_r / n) (ref:vae-hpo.py:130)
        grp = mine[0]
        n = dist.get_world_size(grp)
        flat = torch.arange(1000, dtype=torch.float32) * (rank + 1)
        flat.div_(n)
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=grp)
        res["avg_head"] = flat[:4].tolist()
        res["group_ranks"] = dist.get_process_group_ranks(grp)
    if world_has_all_members(size, ngroups):
        dist.barrier()
    with open(os.path.join(out_dir, f"rank{rank}.json"), "w") as f:
        json.dump(res, f)
    dist.destroy_process_group()


def world_has_all_members(size, ngroups):
    return ngroups > 0 and size % ngroups == 0


if __name__ == "__main__":
    main()
