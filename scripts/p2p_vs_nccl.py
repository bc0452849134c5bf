"""Under torchrun: one DDP group over all ranks, a few training steps with the peer-memory gradient exchange and again
with NCCL from the same initial state and batches; the parameters must agree to fp32 rounding (different summation
order) and be bit-identical across the ranks of the group in both runs.  Prints one line on rank 0."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from multidisttorch_b200.engine import VAEEngine  # noqa: E402
from oracle import vae_numpy as O  # noqa: E402

H = int(os.environ.get("H", "400"))
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
group = dist.new_group(ranks=list(range(world)))
B, steps = 64, 6
xs = [torch.rand(B, 784, generator=torch.Generator().manual_seed(100 * rank + t)).cuda() for t in range(steps)]
out = {}
for mode in ("force", "0"):
    os.environ["MDT_P2P_ALLREDUCE"] = mode
    eng = VAEEngine(hidden=H, max_batch=B, group=group)
    eng.load_segments(O.init_params(5, O.Dims(H=H)))
    eng.set_rng(77 + rank, 0)
    eng.broadcast_params()
    losses = [float(eng.train_step(x).item()) for x in xs]
    en, to = eng.exchange_status()
    assert en == (mode == "force") and not to, (mode, en, to)
    flat = eng.params.clone()
    gathered = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(gathered, flat, group=group)
    assert all(torch.equal(gathered[0], g) for g in gathered), f"ranks diverged ({mode})"
    out[mode] = (flat, losses)
    eng.close()
a, b = out["force"][0], out["0"][0]
rel = float((a - b).norm() / b.norm())
assert rel < 2e-6, rel
assert all(abs(x - y) <= 1e-5 * abs(y) for x, y in zip(out["force"][1], out["0"][1]))
if rank == 0:
    print(f"p2p_vs_nccl world={world} H={H}: params rel diff {rel:.2e}, ranks bit-identical, losses agree")
dist.destroy_process_group()
