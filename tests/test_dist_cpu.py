"""The reference's own multi-node technique, automated: ps + worker OS processes on localhost
(README.md:11-16), CPU math, gloo transport — BASELINE.json config 1 (P1, P3, P4, P7-P10, P13, P17, P30)."""
import os
import re
import subprocess
import sys
import time

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ENV = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""), CUDA_VISIBLE_DEVICES="")
PROGRESS = re.compile(r"Step: (\d+),  Global Step: +(\d+),  Epoch: +(\d+),  Batch: +(\d+) of +(\d+),  Cost: ([\d.naninf]+),"
                      r"  AvgTime: +([\d.]+)ms")


def launch(tmp_path, args, timeout=170):
    r = subprocess.run([sys.executable, "-m", "b200ps.launch", "--workdir", str(tmp_path), "--timeout", str(timeout - 20)] + args,
                       env=ENV, capture_output=True, text=True, timeout=timeout)
    return r.returncode, r.stdout + r.stderr


def lines_of(out, tag):
    return [l[len(tag) + 3:] for l in out.splitlines() if l.startswith(f"[{tag}] ")]


EXAMPLE = ["--backend", "gloo", "--train_examples", "1500", "--test_examples", "300", "--training_epochs", "2",
           "--learning_rate", "0.01", "--frequency", "5"]


def test_example_async_one_ps_two_workers(tmp_path):
    rc, out = launch(tmp_path, ["--num_ps", "1", "--num_workers", "2", "example", "--"] + EXAMPLE)
    assert rc == 0, out
    for w in ("worker:0", "worker:1"):
        ls = lines_of(out, w)
        assert ls[0] == f"worker {w[-1]} start ..." and "Variables initialized ..." in ls and ls[-1] == "done"
        prog = [PROGRESS.fullmatch(l) for l in ls if l.startswith("Step:")]
        assert prog and all(prog), [l for l in ls if l.startswith("Step:")]
        assert any(l.startswith("Test-Accuracy: ") for l in ls) and any(l.startswith("Total Time: ") for l in ls)
        assert any(re.fullmatch(r"Final Cost: [\d.]+", l) for l in ls)
    # async: global_step counts every worker's update => ends at 2 workers x 2 epochs x 15 batches
    finals = [int(PROGRESS.fullmatch([l for l in lines_of(out, w) if l.startswith("Step:")][-1]).group(2))
              for w in ("worker:0", "worker:1")]
    # (the first finisher has at least its own 30 in; how many of the other's depends on when that one started)
    assert max(finals) == 60 and min(finals) >= 30
    ps = lines_of(out, "ps:0")
    assert ps[0] == "ps 0 start ..." and any("Terminating parameter server: 0" in l for l in ps)
    # per-task summary directories with readable event files (example.py:48-51,154,171)
    from b200ps.io import events
    for d in ("ps_0", "worker_0", "worker_1"):
        assert os.path.isdir(tmp_path / "logs" / "mnist" / d)
    recs = list(events.read_events(events.find_event_files(str(tmp_path / "logs" / "mnist" / "worker_1"))[0]))
    assert len(recs) == 1 + 30 and set(recs[5][2]) == {"cost", "accuracy"}


def test_example_sync_two_ps_counts_one_step_per_aggregate(tmp_path):
    rc, out = launch(tmp_path, ["--num_ps", "2", "--num_workers", "2", "example", "--", "--sync"] + EXAMPLE)
    assert rc == 0, out
    finals = []
    for w in ("worker:0", "worker:1"):
        m = PROGRESS.fullmatch([l for l in lines_of(out, w) if l.startswith("Step:")][-1])
        finals.append((int(m.group(2)), float(m.group(6))))
    assert finals[0][0] == finals[1][0] == 30          # 30 aggregated updates, not 60
    assert all(np.isfinite(c) for _, c in finals)
    assert sum("Terminating parameter server" in l for l in out.splitlines()) == 2


def _spawn(tmp_path, job, idx, extra, ports):
    cmd = [sys.executable, "-m", "b200ps.apps.example", f"--job_name={job}", f"--task_index={idx}",
           "--ps_hosts", f"localhost:{ports[0]}", "--worker_hosts", f"localhost:{ports[1]},localhost:{ports[2]}"] + extra
    return subprocess.Popen(cmd, cwd=tmp_path, env=ENV, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)


def test_late_chief_others_wait_for_initialisation(tmp_path):
    from b200ps.launch import free_ports
    ports = free_ports(3)
    extra = ["--backend", "gloo", "--train_examples", "500", "--test_examples", "100", "--training_epochs", "1"]
    ps = _spawn(tmp_path, "ps", 0, extra, ports)
    w1 = _spawn(tmp_path, "worker", 1, extra, ports)
    time.sleep(3)                                    # the chief shows up late
    w0 = _spawn(tmp_path, "worker", 0, extra, ports)
    outs = [p.communicate(timeout=150)[0] for p in (w0, w1, ps)]
    assert w0.returncode == 0 and w1.returncode == 0 and ps.returncode == 0, outs
    assert "chief: running init_op" in outs[0]
    assert "variables initialised by the chief" in outs[1] and "done" in outs[1]


def test_checkpoint_save_and_restore_through_supervisor(tmp_path):
    logdir = str(tmp_path / "ckpt")
    args = ["--num_ps", "1", "--num_workers", "1", "example", "--"] + EXAMPLE + ["--logdir", logdir, "--training_epochs", "1"]
    rc, out = launch(tmp_path, args)
    assert rc == 0, out
    from b200ps.io import saver
    latest = saver.latest_checkpoint(logdir)
    assert latest and latest.endswith("model.ckpt-15")
    v = saver.Saver.restore(latest)
    assert v["weights/Variable"].shape == (784, 100) and v["weights/Variable_1"].shape == (100, 10)
    assert v["biases/Variable"].shape == (100,) and v["biases/Variable_1"].shape == (10,)
    assert v["global_step"].shape == () and v["global_step"].dtype == np.float32 and v["global_step"] == 15
    rc, out = launch(tmp_path, args)                 # second run resumes: the step counter continues
    assert rc == 0 and "chief: restoring from" in out, out
    assert saver.latest_checkpoint(logdir).endswith("model.ckpt-30")


def test_fault_injection_sync_times_out_async_survives(tmp_path):
    base = ["--num_ps", "1", "--num_workers", "2"]
    fault = ["--fault_inject_step", "3", "--fault_rank", "1", "--barrier_timeout_secs", "8"]
    rc, out = launch(tmp_path / "s", base + ["example", "--", "--sync"] + EXAMPLE + fault, timeout=120)
    assert rc != 0 and "fault injection" in out
    assert "done" not in lines_of(out, "worker:0")            # the survivor cannot make progress in sync mode
    rc, out = launch(tmp_path / "a", base + ["--keep_going", "example", "--"] + EXAMPLE + fault, timeout=120)
    assert "fault injection" in out and lines_of(out, "worker:0")[-1] == "done"   # async: survivor finishes


def test_lr2_app_queue_mode_and_log_lines(tmp_path):
    rc, out = launch(tmp_path, ["--num_ps", "1", "--num_workers", "1", "lr2", "--", "--backend", "gloo", "--train", "synthetic",
                                "--test", "synthetic", "--features", "3000", "--num_epochs", "6", "--batch_size", "100",
                                "--learning_rate", "0.5", "--synthetic_rows", "400", "--mode", "queue",
                                "--trace_step_interval", "2"])
    assert rc == 0, out
    w = lines_of(out, "worker:0")
    trace = [l for l in w if re.search(r"\[worker:0\] epoch: \d+, local step: \d+, global step: \d+, loss: [\d.]+", l)]
    assert len(trace) >= 6                                       # one per "epoch" (= step in queue mode) + interval traces
    assert any(re.search(r"Finish evaluate, auc: \(0\.\d+, 0\.\d+\)", l) for l in w)
    assert re.search(r"\[\d{4}-\d\d-\d\d \d\d:\d\d:\d\d\] \[worker:0\] Start train \.\.\.", out)


def test_lr2_native_loader_on_the_gloo_backend(tmp_path):
    """--native_loader on the CPU backend: the C++ reader threads parse, sample and batch this worker's files and the
    CSR arrays go straight into the CPU step (no per-line Python, lr2.py:57-84's hot spot)."""
    rc, out = launch(tmp_path, ["--num_ps", "1", "--num_workers", "2", "lr2", "--", "--backend", "gloo", "--train", "synthetic",
                                "--test", "synthetic", "--features", "3000", "--num_epochs", "3", "--batch_size", "100",
                                "--learning_rate", "0.5", "--synthetic_rows", "400", "--native_loader",
                                "--trace_step_interval", "2"])
    assert rc == 0, out
    for w in ("worker:0", "worker:1"):
        assert any("native loader: " in l and "'malformed': 0" in l and "'batches': 4" in l for l in lines_of(out, w)), out
    ls = lines_of(out, "worker:0")               # the one synthetic test file is worker 0's share (lr2.py:232-253)
    assert any(re.search(r"epoch: \d+, local step: \d+, global step: \d+, loss: [\d.]+", l) for l in ls), out
    assert any(re.search(r"Finish evaluate, auc: \(0\.\d+, 0\.\d+\)", l) for l in ls), out


def test_lr2_round_robin_vs_flat_shard_two_ps(tmp_path):
    for placement in ("round_robin_variable", "flat_shard"):
        rc, out = launch(tmp_path / placement, ["--num_ps", "2", "--num_workers", "2", "lr2", "--", "--backend", "gloo",
                                                "--train", "synthetic", "--test", "synthetic", "--features", "2000",
                                                "--num_epochs", "2", "--batch_size", "200", "--synthetic_rows", "400",
                                                "--placement", placement, "--learning_rate", "0.2"])
        assert rc == 0, out
        assert "Finish train." in out


def test_lr2_wide_deep_over_gloo_and_debug_driver(tmp_path):
    """--deep_dim adds the [F,128] embedding + [128,1] layer of lr2_debug.py:423-428; on the gloo backend the
    rows of both tables travel as messages.  The debug driver counts the lines of its shard of a file list."""
    rc, out = launch(tmp_path / "wd", ["--num_ps", "2", "--num_workers", "2", "lr2", "--", "--backend", "gloo",
                                       "--train", "synthetic", "--test", "synthetic", "--features", "2000",
                                       "--num_epochs", "3", "--batch_size", "200", "--synthetic_rows", "400",
                                       "--deep_dim", "128", "--learning_rate", "0.5"])
    assert rc == 0, out
    assert any(re.search(r"Finish evaluate, auc: \(0\.\d+, 0\.\d+\)", l) for l in lines_of(out, "worker:0")), out
    from b200ps.data.synthetic import write_libsvm_files
    files = write_libsvm_files(str(tmp_path / "data" / "train"), 4, 50, features=500)
    (tmp_path / "data" / "train_file_list").write_text("\n".join(files) + "\n")
    (tmp_path / "cluster_conf.json").write_text('{"ps": ["localhost:2222"], "worker": ["localhost:2223", "localhost:2224"]}')
    for loader in ("thread", "process"):
        r = subprocess.run([sys.executable, "-m", "b200ps.apps.lr2_debug", "--task_index", "1", "--loader", loader,
                            "--test_file_list", ""], cwd=tmp_path, env=ENV, capture_output=True, text=True, timeout=120)
        assert r.returncode == 0, r.stdout + r.stderr
        text = r.stdout + r.stderr
        assert "train lines: 100" in text and "part-00001" in text and "part-00003" in text and "part-00000" not in text


def test_model_export_and_pipeline_demos(tmp_path):
    r = subprocess.run([sys.executable, "-m", "b200ps.apps.model_export", "--device", "cpu", "--work_dir", str(tmp_path / "model")],
                       env=ENV, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    from b200ps.io import exporter
    sigs, variables = exporter.load_export(str(tmp_path / "model" / "00000003"))
    assert sigs == {"inputs": {"x": "Placeholder:0"}, "outputs": {"y": "test/add:0"}}
    assert {"test/weights", "test/bias", "test/test/weights/Adam", "test/test/weights/Adam_1", "test/beta1_power",
            "test/beta2_power"} <= set(variables)
    r = subprocess.run([sys.executable, "-m", "b200ps.apps.input_pipeline_large_dataset", "--pipeline_steps", "120",
                        "--pipeline_print_every", "40"], env=ENV, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "pipeline stopped cleanly" in r.stdout, r.stdout + r.stderr
    assert r.stdout.splitlines()[0].startswith("0 [0.0, 1.0, 2.0")            # FIFO order preserved
    r = subprocess.run([sys.executable, "-m", "b200ps.apps.input_pipeline"], env=ENV, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.count("[") == 30, r.stdout + r.stderr   # 20 train + 10 test label batches
    r = subprocess.run([sys.executable, "-m", "b200ps.apps.hdfs_test"], env=ENV, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "total lines: 300" in r.stdout


def test_shell_launchers(tmp_path):
    from b200ps.data.synthetic import write_libsvm_files
    write_libsvm_files(str(tmp_path / "data" / "train"), 3, 200, features=500)
    write_libsvm_files(str(tmp_path / "data" / "test"), 1, 200, features=500)
    r = subprocess.run(["bash", os.path.join(ROOT, "scripts", "run_lr2.sh"), "--help"], capture_output=True, text=True)
    assert r.returncode == 0 and "--load_mode" in r.stderr and "-j,--job_name" in r.stderr
    r = subprocess.run(["bash", os.path.join(ROOT, "scripts", "run.sh"), "-t", "data/train", "-T", "data/test"], cwd=tmp_path,
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert len(open(tmp_path / "data" / "train_file_list").read().split()) == 3
    r = subprocess.run(["bash", os.path.join(ROOT, "scripts", "run_lr2.sh"), "--run_mode", "product", "--num_ps", "1",
                        "--num_workers", "1", "-t", "data/train", "-T", "data/test", "--load_mode", "all",
                        "--num_epochs", "1", "--features", "500", "--output", "out", "--backend", "gloo"],
                       cwd=tmp_path, env=ENV, capture_output=True, text=True, timeout=150)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "train data: data/train/part-00000,data/train/part-00001,data/train/part-00002" in r.stdout
    assert "Finish evaluate, auc:" in r.stdout


def test_control_plane_liveness_check_detects_a_dead_rank(tmp_path):
    """Context.check_peers over gloo: all ranks answer -> returns; one has died -> PeerLost on the survivor within the
    timeout; Context.abandon leaves without barriers and the interpreter exits normally (the GPU twin of this test is
    tests/test_gpu_multi.py::test_nvlink_backend_detects_a_dead_rank)."""
    import json
    out = tmp_path / "liveness.json"
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
                        "127.0.0.1", "--master-port", "29711", os.path.join(os.path.dirname(__file__), "cpu_liveness_worker.py"),
                        str(out)], env=ENV, capture_output=True, text=True, timeout=180)
    assert r.returncode == 0, (r.stdout + r.stderr)[-3000:]
    res = json.load(open(out))
    assert res.get("lost") and res["checks_ok"] == 3 and res["seconds"] < 10.0 and "stopped answering" in res["message"], res
