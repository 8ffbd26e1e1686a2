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


def _bashflags(script_body, *args):
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script = ". %s/scripts/bashflags\n%s" % (root, script_body)
    return subprocess.run(["bash", "-c", script, "prog"] + list(args), capture_output=True, text=True)


def test_bashflags_types_spellings_and_errors():
    """scripts/bashflags: the shflags role (typed DEFINE_*, long/short/=/--no, multi flags, help, validation)."""
    body = """
DEFINE_string job_name ps "job" j
DEFINE_integer task_index 0 "idx" i
DEFINE_float lr 0.001 "step"
DEFINE_boolean dry false "dry run" n
DEFINE_boolean verbose true "talk" v
DEFINE_multi_string extra "" "repeatable" x
FLAGS "$@" || exit $?
eval set -- "${FLAGS_ARGV}"
echo "job=${FLAGS_job_name} idx=${FLAGS_task_index} lr=${FLAGS_lr} dry=${FLAGS_dry} verbose=${FLAGS_verbose} n=${#FLAGS_extra[@]} last=${FLAGS_extra[1]:-} argc=$# args=$*"
"""
    r = _bashflags(body, "-j", "worker", "-i3", "--lr=1e-2", "-nv", "--noverbose", "-x", "a", "--extra", "b c", "pos 1", "-7", "--", "--raw")
    assert r.returncode == 0, r.stderr
    assert r.stdout.strip() == "job=worker idx=3 lr=1e-2 dry=0 verbose=1 n=2 last=b c argc=3 args=pos 1 -7 --raw"
    r = _bashflags(body)
    assert r.stdout.strip() == "job=ps idx=0 lr=0.001 dry=1 verbose=0 n=0 last= argc=0 args="
    for bad, msg in ((["--task_index", "abc"], "invalid integer"), (["--lr", "fast"], "invalid float"),
                     (["--bogus"], "unknown flag --bogus"), (["-q"], "unknown flag -q"), (["--job_name"], "needs a value"),
                     (["--dry=maybe"], "invalid boolean")):
        r = _bashflags(body, *bad)
        assert r.returncode == 2 and msg in r.stderr, (bad, r.stderr)
    r = _bashflags(body, "--help")
    assert r.returncode == 1 and "-j,--job_name:  job (string; default: 'ps')" in r.stderr and "--[no]dry" in r.stderr
    r = _bashflags(body, "--helpxml")
    assert "<name>task_index</name>" in r.stdout and "<type>integer</type>" in r.stdout


def test_bashflags_definition_errors_required_and_passthrough():
    r = _bashflags('DEFINE_integer n abc "bad default"; echo rc=$?')
    assert "rc=2" in r.stdout and "invalid integer" in r.stderr
    r = _bashflags('DEFINE_string a x "1" s; DEFINE_string b y "2" s; echo rc=$?')
    assert "rc=2" in r.stdout and "defined twice" in r.stderr
    r = _bashflags('DEFINE_string ARGV x "reserved"; echo rc=$?')
    assert "rc=2" in r.stdout and "reserved" in r.stderr
    r = _bashflags('DEFINE_string train "" "t"; flags_required train; FLAGS "$@"; echo rc=$?')
    assert "rc=2" in r.stdout and "required flag(s) missing: --train" in r.stderr
    r = _bashflags('FLAGS_PASSTHROUGH=1; DEFINE_integer k 1 "k"; FLAGS "$@" || exit 9; echo "k=${FLAGS_k} rest=${FLAGS_ARGV}"',
                   "--k", "5", "--backend", "gloo", "--sync")
    assert r.stdout.strip() == "k=5 rest=--backend gloo --sync"
    r = _bashflags('DEFINE_integer k 1 "k"; flags_reset; DEFINE_string k z "again"; FLAGS "$@"; echo "k=${FLAGS_k}"', "--k", "w")
    assert r.stdout.strip() == "k=w"


def test_trace_helpers_step_timer_and_disabled_nvtx():
    import time
    from b200ps.utils import StepTimer, nvtx_range
    with nvtx_range("noop"):            # B200PS_NVTX unset: a plain pass-through, no CUDA needed
        x = 1
    t = StepTimer()
    time.sleep(0.02); t.tick(4)
    avg = t.report()
    assert x == 1 and 3.0 < avg < 50.0 and t.report() >= 0.0


def test_cpu_thread_cap_for_single_host_clusters(monkeypatch):
    """Tasks started by hand on one host share its cores: every task keeps cores / (2 x tasks) intra-op threads,
    unless OMP_NUM_THREADS is set or the cluster spans other hosts (then the choice is the user's)."""
    import os
    import torch
    from b200ps.cluster.server import _cap_cpu_threads
    from b200ps.cluster.spec import ClusterSpec
    before = torch.get_num_threads()
    try:
        torch.set_num_threads(max(before, 2))
        monkeypatch.setenv("OMP_NUM_THREADS", "7")
        _cap_cpu_threads(ClusterSpec.local(1, 2))
        assert torch.get_num_threads() == max(before, 2)                      # the user chose: untouched
        monkeypatch.delenv("OMP_NUM_THREADS")
        _cap_cpu_threads(ClusterSpec({"ps": ["10.0.0.1:2222"], "worker": ["localhost:2223"]}))
        assert torch.get_num_threads() == max(before, 2)                      # other hosts: untouched
        _cap_cpu_threads(ClusterSpec.local(1, 2))
        want = max(1, min(max(before, 2), (os.cpu_count() or 1) // 6))
        assert torch.get_num_threads() == want
    finally:
        torch.set_num_threads(before)
