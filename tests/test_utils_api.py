"""Kept host API (utils.py mirror) against the known answers the reference
itself produced (tests/golden/api.json, written by oracle/make_golden.py).
CPU / gloo only."""
import json
import os

import pytest

from multidisttorch_b200 import utils as U
from tests import dist_launch
from tests.golden_cases import GOLD, load

API = json.load(open(os.path.join(GOLD, "api.json")))


def test_parse_slurm_nodelist_known_answers():
    for s, exp in API["parse_slurm_nodelist"].items():
        assert U.parse_slurm_nodelist(s) == exp, s
    with pytest.raises(AttributeError):      # reference: re.match(...) is None -> .groups() raises
        U.parse_slurm_nodelist("a[-5]")


def test_find_ifname_loopback():
    assert U.find_ifname("127.0.0.1") == API["find_ifname_loopback"]
    assert U.find_ifname("127.0.0.1") is not None


def test_init_comm_size_and_rank_env_precedence(monkeypatch):
    keys = ["OMPI_COMM_WORLD_SIZE", "OMPI_COMM_WORLD_RANK", "SLURM_NPROCS", "SLURM_PROCID", "RANK", "WORLD_SIZE"]
    for case in API["init_comm_size_and_rank"]:
        for k in keys:
            monkeypatch.delenv(k, raising=False)
        for k, v in case["env"].items():
            monkeypatch.setenv(k, v)
        assert list(U.init_comm_size_and_rank()) == case["out"], case["env"]


def test_get_comm_without_init():
    assert U.get_comm_size_and_rank() == (1, 0)


def test_star_export_names_equal_reference():
    ns = {}
    exec("from utils import *", ns)
    assert sorted(k for k in ns if not k.startswith("__")) == API["star_names"]


def test_group_ranks_rule():
    assert U.group_ranks(8, 8) == [[i] for i in range(8)]
    assert U.group_ranks(8, 4) == [[0, 1], [2, 3], [4, 5], [6, 7]]
    assert U.group_ranks(8, 2) == [[0, 1, 2, 3], [4, 5, 6, 7]]
    assert U.group_ranks(8, 1) == [list(range(8))]
    assert U.group_ranks(8, 3) == [[0, 1], [2, 3], [4, 5]]          # 6, 7 orphaned
    assert U.group_ranks(7, 2) == [[0, 1, 2], [3, 4, 5]]            # 6 orphaned


@pytest.mark.parametrize("world,ngroups", [(2, 2), (4, 2), (5, 2)])
def test_setup_ddp_groups_membership_and_prints_bit_exact(world, ngroups):
    gold = next(g for g in API["groups"] if g["world"] == world and g["ngroups"] == ngroups)
    out, codes, logs = dist_launch.launch(world, ["plain", ngroups, "{out}"])
    assert codes == [0] * world, logs
    res = dist_launch.results(out, world)
    assert [r["membership"] for r in res] == gold["membership"]
    # the master port differs between the golden run and this one; everything else is byte-identical
    norm = lambda s: "\n".join(l.rsplit(":", 1)[0] if l.startswith("Distributed data parallel") else l
                               for l in s.splitlines())
    assert norm(res[0]["stdout"]) == norm(gold["stdout_rank0"])
    assert norm(res[-1]["stdout"]) == norm(gold["stdout_last"])
    assert all(r["get_comm"] == [world, r["rank"]] for r in res)


def test_subgroup_all_gather_and_print0_like_example_subgroup():
    world, ngroups = 4, 2
    out, codes, logs = dist_launch.launch(world, ["gather", ngroups, "{out}"])
    assert codes == [0] * world, logs
    res = dist_launch.results(out, world)
    assert [r["gather"] for r in res] == [[0, 1], [0, 1], [2, 3], [2, 3]]
    gold = load("ddp_2x2_b16")
    for r in range(world):
        assert res[r]["print0"] == str(gold[f"print0.rank{r}"])


def test_group_gradient_average_over_gloo():
    world, ngroups = 4, 2
    out, codes, logs = dist_launch.launch(world, ["allreduce", ngroups, "{out}"])
    assert codes == [0] * world, logs
    res = dist_launch.results(out, world)
    # group 0 = ranks 0,1 -> mean factor (1+2)/2 ; group 1 = ranks 2,3 -> (3+4)/2
    for r, f in zip(range(4), (1.5, 1.5, 3.5, 3.5)):
        assert res[r]["avg_head"] == [0.0, f, 2 * f, 3 * f]
        assert res[r]["group_ranks"] == ([0, 1] if r < 2 else [2, 3])
