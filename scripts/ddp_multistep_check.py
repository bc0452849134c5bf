"""A DDP group stepped through mdt_train_steps (16-step graphs: staged-ahead batches, weight-shadow refresh on a side
stream) against the same group stepped one mdt_train_step at a time: parameters must be bit-identical, on every rank and
across ranks, for every exchange.  torchrun --nproc-per-node N scripts/ddp_multistep_check.py

The two engines are stepped ONE AFTER THE OTHER (synchronize + barrier in between): the peer-memory exchange kernels
spin with every block resident, so two of them in flight on the same GPUs in different orders on different ranks
starve each other's GEMMs until the bounded wait (30 s) expires and flags the step.  The first version of this script
did not do that; its one 2-GPU run (round 2) reported False for tf32x3/p2p (tf32x3/nccl: True) and overran its time
limit -- an expired wait, from that overlap or from a real fault of the multi-step peer-memory path: not known
(DESIGN.md section 5).  bench.py --gpus N runs the same comparison before it times a layout (multistep_parity)."""

import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from multidisttorch_b200 import synth
from multidisttorch_b200.engine import VAEEngine
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("cpu:gloo,cuda:nccl", device_id=torch.device("cuda", local))
g = dist.new_group(ranks=list(range(world)))
pool = torch.rand(5, 64, 784, generator=torch.Generator().manual_seed(100 + rank)).cuda()
ok_all = True
for mode in ("tf32x3", "bf16"):
    for xch in ("p2p", "nccl", "p2p_two_shot"):
        engs = []
        for _ in range(2):
            e = VAEEngine(max_batch=64, gemm_mode=mode, group=g, exchange=xch)
            e.load_segments(synth.init_params(9)); e.broadcast_params(); e.set_rng(50 + rank, 0)
            engs.append(e)
        a, b = engs
        a.train_steps(pool, 2, 37)
        a.synchronize(); dist.barrier(group=g)
        for i in range(37):
            b.train_step(pool[(2 + i) % 5], None, want_loss=False)
        a.synchronize(); b.synchronize()
        same = torch.equal(a.params, b.params) and torch.equal(a.exp_avg_sq, b.exp_avg_sq)
        got = [torch.empty_like(a.params) for _ in range(world)]
        dist.all_gather(got, a.params.clone(), group=g)
        same = same and all(torch.equal(got[0], t) for t in got) and not a.exchange_status()[1]
        flag = torch.tensor([1 if same else 0], device="cuda"); dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        ok_all = ok_all and bool(flag.item())
        if rank == 0:
            print(f"[ddp_multistep_check] {mode}/{xch}: multi-step == single-step, ranks identical: {bool(flag.item())}", flush=True)
        a.close(); b.close()
dist.barrier(); dist.destroy_process_group()
sys.exit(0 if ok_all else 1)
