"""One DDP group over every visible GPU at H=4096 through the NCCL exchange (DDP's two buckets), stepped with
mdt_train_steps: the case whose 16-step capture failed on 8 ranks.  torchrun --nproc-per-node N scripts/nccl_h4096_check.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from multidisttorch_b200 import synth
from multidisttorch_b200.engine import VAEEngine
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("cpu:gloo,cuda:nccl", device_id=torch.device("cuda", local))
g = dist.new_group(ranks=list(range(world)))
pool = torch.rand(4, 128, 784, device="cuda")
for xch in ("nccl", "auto"):
    e = VAEEngine(hidden=4096, max_batch=128, group=g, exchange=xch)
    e.load_segments(synth.init_params(5, H=4096)); e.broadcast_params(); e.set_rng(3 + rank, 0)
    losses = torch.zeros(40, device="cuda")
    e.train_steps(pool, 0, 40, losses)
    e.synchronize()
    flat = e.params.clone(); got = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(got, flat, group=g)
    ok = all(torch.equal(got[0], t) for t in got) and bool(torch.isfinite(losses).all())
    if rank == 0:
        print(f"[nccl_h4096_check] exchange={xch} p2p={e.p2p}: 40 steps, ranks identical and finite: {ok}; loss {losses[0].item():.1f} -> {losses[-1].item():.1f}", flush=True)
    e.close()
dist.barrier(); dist.destroy_process_group()
