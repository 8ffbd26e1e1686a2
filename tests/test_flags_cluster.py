"""Flags, ClusterSpec, placement (P1, P2, P5, P6) — CPU only."""
import json

import pytest

from b200ps import flags as F
from b200ps.cluster.spec import ClusterSpec
from b200ps.ps.placement import VarSpec, plan_layout


def test_reference_flag_names_and_defaults():
    import b200ps.apps.lr2  # noqa: F401  (defines the lr2 flags)
    F.FLAGS._reset()
    F.parse([])
    d = F.FLAGS.as_dict()
    # lr2.py:20-35
    assert (d["job_name"], d["task_index"], d["thread_num"]) == ("worker", 0, 2)
    assert d["learning_rate"] == 0.001 and d["num_epochs"] == 120 and d["batch_size"] == 500
    assert d["features"] == 4762348 and d["trace_step_interval"] == 10000
    assert d["train_sampling_rate"] == 1.0 and d["test_sampling_rate"] == 1.0 and d["mode"] == "all"
    assert d["train_queue_capacity"] == 2000 and d["test_queue_capacity"] == 2000
    assert d["train"].startswith("hdfs://") and d["checkpoint"].startswith("hdfs://")


def test_flag_parsing_forms_and_set_without_define():
    F.DEFINE_integer("t_int", 3, "x")
    F.DEFINE_boolean("t_bool", False, "x")
    F.DEFINE_string("t_str", "a", "x", short="q")
    rest = F.parse(["--t_int=7", "--t_bool", "-q", "zz", "--unknown_flag", "1"])
    assert F.FLAGS.t_int == 7 and F.FLAGS.t_bool is True and F.FLAGS.t_str == "zz"
    assert rest == ["--unknown_flag", "1"]
    F.parse(["--not_bool"])                   # gflags/shflags style negation
    assert F.FLAGS.t_bool is False
    F.FLAGS.work_dir = "./model/"            # model_export.py:10 sets undeclared flags
    assert F.FLAGS.work_dir == "./model/"
    with pytest.raises(AttributeError):
        F.FLAGS.never_defined


def test_cluster_spec_json_and_ranks(tmp_path):
    p = tmp_path / "cluster_conf.json"
    p.write_text(json.dumps({"ps": ["localhost:2222"], "worker": ["localhost:2223", "localhost:2224"]}))
    c = ClusterSpec.from_json(str(p))
    assert c.num_ps == 1 and c.num_workers == 2 and c.world_size() == 3
    assert c.task_address("worker", 1) == "localhost:2224"
    assert c.global_rank("worker", 1) == 1 and c.global_rank("ps", 0) == 2
    assert c.role_of_rank(2) == ("ps", 0) and c.role_of_rank(0) == ("worker", 0)
    assert c.rendezvous() == ("127.0.0.1", 2222)
    with pytest.raises(IndexError):
        c.task_address("worker", 2)
    with pytest.raises(ValueError):
        ClusterSpec({"ps": ["nocolon"]})
    assert ClusterSpec({"worker": {1: "b:2", 0: "a:1"}}).job_tasks("worker") == ["a:1", "b:2"]
    # the repository ships the reference's cluster file verbatim
    shipped = ClusterSpec.from_json("cluster_conf.json")
    assert shipped.as_dict() == {"ps": ["localhost:2222"], "worker": ["localhost:2223"]}


def test_round_robin_placement_matches_replica_device_setter():
    # creation order global_step, W1, W2, b1, b2 over 2 ps -> ps0, ps1, ps0, ps1, ps0
    order = {"global_step": 0, "weights/Variable": 1, "weights/Variable_1": 2, "biases/Variable": 3,
             "biases/Variable_1": 4}
    specs = [VarSpec("weights/Variable", (784, 100)), VarSpec("weights/Variable_1", (100, 10)),
             VarSpec("biases/Variable", (100,)), VarSpec("biases/Variable_1", (10,))]
    lay = plan_layout(specs, 2, "round_robin_variable", order)
    assert [lay.owners[s.name] for s in specs] == [1, 0, 1, 0]
    # every owner's variables are contiguous and the ranges tile the flat space
    assert lay.bounds[0] == 0 and lay.bounds[-1] == lay.total
    for s in specs:
        o = lay.owners[s.name]
        assert lay.bounds[o] <= lay.offsets[s.name] and lay.offsets[s.name] + s.numel <= lay.bounds[o + 1]
    # lr2: global_step -> ps0, W -> ps1, b -> ps0
    lay2 = plan_layout([VarSpec("bias/Variable", (1,))], 2, "round_robin_variable",
                       {"global_step": 0, "weights/Variable": 1, "bias/Variable": 2})
    assert lay2.owners["bias/Variable"] == 0


def test_flat_shard_placement_is_even_and_aligned():
    specs = [VarSpec("a", (784, 100)), VarSpec("b", (100, 10)), VarSpec("c", (100,)), VarSpec("d", (10,))]
    for world in (1, 2, 4, 8):
        lay = plan_layout(specs, world, "flat_shard")
        sizes = {lay.bounds[r + 1] - lay.bounds[r] for r in range(world)}
        assert len(sizes) == 1 and all(b % 8 == 0 for b in lay.bounds)
        assert lay.total >= sum(s.numel for s in specs)
        assert all(lay.offsets[s.name] % 8 == 0 for s in specs)
    with pytest.raises(ValueError):
        plan_layout(specs, 2, "nope")
