"""One GPU rank of the multi-GPU parity check (launched by tests/test_gpu_multi.py or by hand):
utils.setup_ddp / setup_ddp_groups -> engine on the subgroup's NCCL communicator -> N steps against the golden
vectors the reference's real DDP (gloo) produced."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from utils import *  # noqa: E402,F401,F403
from multidisttorch_b200.engine import VAEEngine  # noqa: E402
from oracle import vae_numpy as O  # noqa: E402
from tests.golden_cases import PARAM_SEED, batch_seed, check_fp, load  # noqa: E402


def main():
    tag, mode, out_dir = sys.argv[1], sys.argv[2], sys.argv[3]
    g = load(tag)
    world, ng, B, n = (int(g[k]) for k in ("world", "ngroups", "B", "n_steps"))
    size, rank = setup_ddp()
    assert size == world, (size, world)
    groups = setup_ddp_groups(ng)
    gid = next(i for i, gr in enumerate(groups) if dist.get_rank(gr) >= 0)
    group = groups[gid]
    grank = dist.get_rank(group)
    xch = os.environ.get("MDT_EXCHANGE") or None            # None: the shipped choice
    eng = VAEEngine(max_batch=B, gemm_mode=mode, group=group, exchange=xch)
    # only group-rank 0 holds the real initial parameters: the broadcast must overwrite the others (DDP ctor)
    eng.load_segments(O.init_params(PARAM_SEED + gid if grank == 0 else 999 + rank))
    eng.broadcast_params()
    # (loss, parameters after the golden's 3-5 steps); the same bounds as bench.py's ddp_parity
    # measured on 2 x B200 (1x2 golden, 5 steps): tf32x3 loss 1.1e-7 / parameters 2.0e-7, bf16 1.7e-5 / 1.4e-2
    tol = {"fp32": (5e-6, 2e-5), "tf32x3": (5e-6, 2e-5), "tf32": (2e-3, 1e-1), "bf16": (6e-4, 1e-1)}[mode]
    # A fixture whose run holds a ReLU pre-activation within fp32 rounding of zero (relu_margin; oracle/make_golden.py
    # run_ddp) cannot pin parameters tightly: one flipped mask moves a weight row by O(lr).  Measured: 2.9e-3 on W1.
    shift = int(g["seed_shift"]) if "seed_shift" in g else 0
    if "relu_margin" in g and float(g["relu_margin"]) < 1.5e-6 and mode in ("fp32", "tf32x3"):
        tol = (tol[0], 1e-2)
    res = {"rank": rank, "gid": gid, "losses": []}
    for t in range(n):
        x, eps = O.make_batch(batch_seed(t, rank) + shift, B)
        loss = float(eng.train_step(torch.from_numpy(x).cuda(), torch.from_numpy(eps).cuda()).item())
        ref = float(g[f"losses.rank{rank}"][t])
        assert abs(loss - ref) <= tol[0] * abs(ref), (tag, mode, rank, t, loss, ref)
        res["losses"].append(loss)
    p = eng.segments_to_numpy()
    for k in O.SEGMENTS:
        check_fp(g, f"g{gid}.paramN", k, p[k], tol[1], f"{tag}/{mode}/rank{rank}")
    # which exchange ran: peer memory on the tensor-core path unless switched off, NCCL otherwise; no expired waits
    n_grp = dist.get_world_size(group)
    want_p2p = (mode != "fp32" and os.environ.get("MDT_P2P_ALLREDUCE", "1") != "0" and xch != "nccl"
                and (n_grp <= int(os.environ.get("MDT_P2P_MAX_RANKS", "4")) or xch in ("p2p", "p2p_two_shot")))
    assert eng.exchange_status() == (want_p2p, False), (eng.exchange_status(), want_p2p)
    # ranks of a group must hold bit-identical parameters
    flat = eng.params.clone()
    gathered = [torch.empty_like(flat) for _ in range(dist.get_world_size(group))]
    dist.all_gather(gathered, flat, group=group)
    assert all(torch.equal(gathered[0], t) for t in gathered), "group ranks diverged"
    res["ok"] = True
    dist.barrier()
    with open(os.path.join(out_dir, f"rank{rank}.json"), "w") as f:
        json.dump(res, f)
    eng.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    try:
        main()
    except BaseException:          # leave at once: peers blocked in a collective are killed by the launcher
        import traceback
        traceback.print_exc()
        sys.stdout.flush(); sys.stderr.flush()
        os._exit(1)
