#!/usr/bin/env python
"""BASELINE.json metric #2: subgroup all-reduce correctness + bus-bandwidth sweep (config 2: 4 groups x 2 GPUs on
8 x B200; any world / group count works).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 bench_allreduce.py --ngroups 4
    python -m torch.distributed.run ... bench_allreduce.py --ngroups 4 --impl reference      # gloo on host cores

Every rank builds ALL groups in order (the rule of ref:utils.py:146-163 / ref:example-subgroup.py:10-23) with the
same rank->group map (`utils.group_ranks`), then all groups all-reduce CONCURRENTLY: in-place fp32 SUM on the
subgroup's NCCL communicator through the C-ABI (`mdt_allreduce_bench`, CUDA-event timed on the device).
busbw = (S / t) * 2(n-1)/n.  Correctness: payload (group_rank + 1) -> every element must equal n(n+1)/2 exactly,
plus the reference-style all_gather of rank ids.  One JSON line on stdout (rank 0)."""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

DDP_SIZES = [1257536, 1353760, 2611296, 26709216]      # DDP bucket / flat-gradient sizes


def sizes(max_bytes):
    s, out = 8, []
    while s <= max_bytes:
        out.append(s)
        s *= 2
    return sorted(set(out + [d for d in DDP_SIZES if d <= max_bytes]))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ngroups", type=int, default=0, help="default: world // 2 groups of 2 ranks")
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--max-bytes", type=int, default=512 << 20)
    ap.add_argument("--iters", type=int, default=20)
    args = ap.parse_args()
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    import torch
    import torch.distributed as dist
    from multidisttorch_b200 import utils as U

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    ngroups = args.ngroups or max(world // 2, 1)
    ref = args.impl == "reference"
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29533")
    if ref:
        dist.init_process_group("gloo", init_method="env://", rank=rank, world_size=world)
        device = torch.device("cpu")
        torch.set_num_threads(max((os.cpu_count() or 1) // world, 1))
    else:
        torch.cuda.set_device(local)
        device = torch.device("cuda", local)
        dist.init_process_group("cpu:gloo,cuda:nccl", init_method="env://", rank=rank, world_size=world, device_id=device)

    ranks = U.group_ranks(world, ngroups)
    groups = [dist.new_group(ranks=r) for r in ranks]            # every rank creates every group, in order
    gid = next((g for g, r in enumerate(ranks) if rank in r), None)
    results = {}
    gather_ok = True
    if gid is not None:
        group = groups[gid]
        n = dist.get_world_size(group)
        grank = dist.get_rank(group)
        # reference-style membership check (ref:example-subgroup.py:25-33)
        got = [torch.tensor([0]) for _ in range(n)]
        dist.all_gather(got, torch.tensor([rank]), group=group)
        gather_ok = [int(t.item()) for t in got] == ranks[gid]
        if not ref:
            from multidisttorch_b200 import _lib
            from multidisttorch_b200.engine import nccl_comm_ptr
            lib = _lib.load()
            comm = C.c_void_p(nccl_comm_ptr(group, device)) if n > 1 else None
            stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        max_bytes = args.max_bytes if not ref else min(args.max_bytes, 64 << 20)
        for S in sizes(max_bytes):
            buf = torch.full((S // 4,), float(grank + 1), dtype=torch.float32, device=device) if S >= 4 else None
            if buf is None:
                buf = torch.full((2,), float(grank + 1), dtype=torch.float32, device=device)
            nbytes = buf.numel() * 4
            # correctness first (one all-reduce), then the timed loop on fresh data
            if ref or n == 1:
                if n > 1:
                    dist.all_reduce(buf, group=group)
            else:
                sec = C.c_double()
                _lib.check(lib.mdt_allreduce_bench(comm, C.c_void_p(buf.data_ptr()), nbytes, 1, stream, C.byref(sec)))
            # mdt_allreduce_bench runs a warm-up + `iters` reductions in place: after k reductions of a constant payload
            # v the value is v0 * n^k with v0 the first sum; check the first reduction through a separate single call
            exp_first = n * (n + 1) / 2
            if ref or n == 1:
                ok = bool((buf == (exp_first if n > 1 else grank + 1)).all())
            else:
                ok = bool((buf == exp_first * n).all())          # warm-up + 1 timed = two reductions
            buf.fill_(1.0)
            dist.barrier(group=group)
            if ref:
                it = max(min(args.iters, int(2e9 // max(nbytes, 1)) + 1), 3)
                t0 = time.perf_counter()
                for _ in range(it):
                    dist.all_reduce(buf, group=group)
                t = (time.perf_counter() - t0) / it
            elif n > 1:
                it = args.iters if nbytes <= (64 << 20) else 5
                buf.fill_(0.0)        # zeros stay zeros: no overflow over many in-place sums
                sec = C.c_double()
                _lib.check(lib.mdt_allreduce_bench(comm, C.c_void_p(buf.data_ptr()), nbytes, it, stream, C.byref(sec)))
                t = sec.value
            else:
                t = float("nan")
            results[nbytes] = {"t": t, "ok": ok}
            del buf
    # collect: per size, time = max over the ranks of a group; report min / median over groups
    gathered = [None] * world
    dist.all_gather_object(gathered, {"gid": gid, "res": results, "gather_ok": gather_ok})
    if rank == 0:
        n = world // ngroups
        table = []
        all_ok = all(g["gather_ok"] for g in gathered)
        for S in sorted(results):
            per_group = []
            for g in range(ngroups):
                ts = [x["res"][S]["t"] for x in gathered if x["gid"] == g and S in x["res"]]
                oks = [x["res"][S]["ok"] for x in gathered if x["gid"] == g and S in x["res"]]
                all_ok = all_ok and all(oks)
                if ts:
                    per_group.append(max(ts))
            if not per_group or n < 2:
                continue
            bw = sorted((S / t) * 2 * (n - 1) / n / 1e9 for t in per_group)
            table.append({"bytes": S, "us": round(max(per_group) * 1e6, 2), "busbw_GBps_min": round(bw[0], 3),
                          "busbw_GBps_median": round(bw[len(bw) // 2], 3)})
        peak = max((r["busbw_GBps_min"] for r in table), default=None)
        out = {"metric": "subgroup allreduce bus GB/s", "impl": args.impl, "unit": "GB/s", "value": peak, "n_gpus": world,
               "layout": f"{ngroups}x{n}", "dtype": "f32", "op": "sum, in place", "correct": all_ok,
               "roof": {"nominal_GBps_per_direction": 900, "measured_reference_GBps": 725,
                        "note": "B200_PROFILING.md: 8-rank all-reduce busbw 725 GB/s at 1 GiB, NVLink 5 nominal 900 GB/s/direction"},
               "ddp_sizes": {str(r["bytes"]): r for r in table if r["bytes"] in DDP_SIZES},
               "sweep": table}
        os.write(real_stdout, (json.dumps(out) + "\n").encode())
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
