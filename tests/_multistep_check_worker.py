"""One gloo rank of tests/test_bench_multistep_check.py: bench.multistep_check driven with fake engines (CPU tensors),
so that its control flow and collectives -- the part of `bench.py --gpus N` that decides whether peer-memory groups are
timed on multi-step graphs -- are exercised without a GPU."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import bench  # noqa: E402


class FakeEngine:
    """params += mean(batch) per step.  No collective inside (a rank whose engine raises must not block its peers'
    fake steps -- on the GPU the peers' steps are asynchronous and end in a bounded wait); the ranks of a group see the
    same pool, so their parameters agree like those of a real group after the gradient exchange."""

    def __init__(self, group, fault=None, p2p=True):
        self.group, self.fault, self.p2p = group, fault, p2p
        self.params = torch.zeros(16)
        self.exp_avg = torch.zeros(16)
        self.exp_avg_sq = torch.zeros(16)
        self.closed = False
        self.timed_out = False

    def _step(self, x):
        g = x.mean().reshape(1).clone()
        self.params += g
        self.exp_avg += 0.1 * g
        self.exp_avg_sq += g * g

    def train_steps(self, pool, first, n):
        if self.fault == "raise_multi":
            raise RuntimeError("injected failure in the multi-step leg")
        for i in range(n):
            self._step(pool[(first + i) % pool.shape[0]])
        if self.fault == "drift":
            self.params[3] += 1e-3
        if self.fault == "timeout":
            self.timed_out = True

    def train_step(self, x, eps, want_loss=True):
        if self.fault == "raise_single":
            raise RuntimeError("injected failure in the single-step leg")
        self._step(x)

    def synchronize(self):
        pass

    def exchange_status(self):
        return self.p2p, self.timed_out

    def close(self):
        self.closed = True


def main():
    out_dir = sys.argv[1]
    rank, world = int(os.environ["OMPI_COMM_WORLD_RANK"]), int(os.environ["OMPI_COMM_WORLD_SIZE"])
    dist.init_process_group("gloo", rank=rank, world_size=world)
    k = 2
    groups = [dist.new_group(ranks=list(range(k * gi, k * gi + k))) for gi in range(world // k)]
    mine = groups[rank // k]
    pool = torch.rand(5, 8, 4, generator=torch.Generator().manual_seed(rank // 2))
    res = {}
    made = []

    def case(name, fault_on=None, fault=None, multi_only=True, **kw):
        count = [0]

        def make():
            count[0] += 1
            f = fault if (fault_on is None or rank == fault_on) and (not multi_only or count[0] == 1) else None
            e = FakeEngine(mine, f, p2p=kw.get("p2p", True))
            made.append(e)
            return e
        res[name] = bench.multistep_check(torch, dist, make, pool, mine, k, "cpu", name, need_p2p=kw.get("need_p2p", False))

    case("identical")
    case("drift_on_rank_1", fault_on=1, fault="drift")                       # one rank's multi-step engine differs
    case("raise_multi_on_rank_0", fault_on=0, fault="raise_multi")           # ... raises: nobody may hang
    case("raise_single_on_rank_1", fault_on=1, fault="raise_single", multi_only=False)
    case("timeout_flag", fault_on=0, fault="timeout")
    case("no_peer_memory", p2p=False, need_p2p=True)
    res["all_closed"] = all(e.closed for e in made)
    with open(os.path.join(out_dir, f"rank{rank}.json"), "w") as f:
        json.dump(res, f)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
