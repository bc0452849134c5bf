"""bench.multistep_check (the multi-step-graph vs single-step comparison `bench.py --gpus N` runs on every layout's
groups before timing them) under gloo with fake engines: same verdict on every rank, no hang when one rank's engine
raises, engines always closed."""
import os

from tests import dist_launch

HERE = os.path.dirname(os.path.abspath(__file__))


def test_multistep_check_verdicts_agree_across_ranks_and_failures_do_not_hang():
    world = 4          # two groups of two: a fault in one group must reach the other group's ranks too
    out, codes, logs = dist_launch.launch(world, ["{out}"], script=os.path.join(HERE, "_multistep_check_worker.py"), timeout=120)
    assert codes == [0] * world, "\n".join(logs)
    res = dist_launch.results(out, world)
    for r in res:
        assert r["identical"] == "ok"
        assert r["drift_on_rank_1"] == "FAIL"
        assert r["raise_multi_on_rank_0"] == "FAIL"
        assert r["raise_single_on_rank_1"] == "FAIL"
        assert r["timeout_flag"] == "FAIL"
        assert r["no_peer_memory"] is None
        assert r["all_closed"] is True
    assert any("injected failure in the multi-step leg" in log for log in logs)
