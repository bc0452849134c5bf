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


def _scripted(results):
    """check_kinds stand-in: pops the next scripted verdict table; records what was asked."""
    asked = []

    def check(kinds, tag=""):
        asked.append((tuple(kinds), tag))
        table = results.pop(0)
        return {k: table[k] for k in kinds if table.get(k) is not None}
    return check, asked


def test_multistep_fallbacks_all_ok_sets_nothing():
    import bench
    env = {}
    check, asked = _scripted([{"p2p": "ok", "p2p_two_shot": "ok", "nccl": "ok"}])
    chk = bench.multistep_fallbacks(check, env)
    assert chk == {"p2p": "ok", "p2p_two_shot": "ok", "nccl": "ok"} and env == {} and len(asked) == 1


def test_multistep_fallbacks_first_step_back_is_the_shadow_refresh():
    import bench
    env = {}
    check, asked = _scripted([{"p2p": "FAIL", "p2p_two_shot": "ok", "nccl": "ok"}, {"p2p": "ok"}])
    chk = bench.multistep_fallbacks(check, env)
    assert env == {"MDT_NO_DEFER_SHADOWS": "1"}
    assert asked[1] == (("p2p",), "/no_defer") and chk["retry_with_MDT_NO_DEFER_SHADOWS"] == {"p2p": "ok"}
    assert "MDT_NO_DEFER_SHADOWS=1" in chk["note"] and "SINGLE_STEPS" not in chk["note"]


def test_multistep_fallbacks_peer_memory_ends_on_single_steps_nccl_does_not():
    import bench
    env = {}
    check, _ = _scripted([{"p2p": "FAIL", "p2p_two_shot": None, "nccl": {"error": "boom"}}, {"p2p": "FAIL", "nccl": "FAIL"}])
    chk = bench.multistep_fallbacks(check, env)
    assert env == {"MDT_NO_DEFER_SHADOWS": "1", "MDT_P2P_SINGLE_STEPS": "1"}
    assert "MDT_P2P_SINGLE_STEPS=1" in chk["note"]
    # a later layout of the same run: the switches are already set, nothing is retried
    check2, asked2 = _scripted([{"p2p": "ok", "p2p_two_shot": "ok", "nccl": "FAIL"}])
    chk2 = bench.multistep_fallbacks(check2, env)
    assert len(asked2) == 1 and "note" not in chk2
