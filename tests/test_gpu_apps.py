"""The reference's command lines on the NVLink backend, one worker on one GPU (rank 0 = worker 0 + ps shard 0):
example.py's MNIST trainer and lr2.py's sparse trainer through `b200ps.launch`, as a user would start them
(P1, P11-P13, P15-P17, P30).  The multi-GPU variants are in test_gpu_multi.py."""
import os
import re
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ENV = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))


def launch(tmp_path, args, timeout=280):
    r = subprocess.run([sys.executable, "-m", "b200ps.launch", "--workdir", str(tmp_path), "--timeout", str(timeout - 20),
                        "--num_ps", "0", "--num_workers", "1"] + args, env=ENV, capture_output=True, text=True, timeout=timeout)
    return r.returncode, r.stdout + r.stderr


def test_example_app_on_the_nvlink_backend(native_lib, tmp_path):
    rc, out = launch(tmp_path, ["example", "--", "--backend", "nvlink", "--train_examples", "1500", "--test_examples", "300",
                                "--training_epochs", "2", "--learning_rate", "0.01", "--frequency", "5"])
    assert rc == 0, out[-3000:]
    steps = re.findall(r"Step: (\d+),  Global Step: +(\d+),  Epoch: +(\d+),  Batch: +(\d+) of +(\d+),  Cost: ([\d.]+)", out)
    assert steps and int(steps[-1][1]) == 30 and steps[-1][2] == "2"           # 2 epochs x 15 batches, one worker
    assert re.search(r"Test-Accuracy: [\d.]+", out) and re.search(r"Final Cost: [\d.]+", out) and "done" in out


@pytest.mark.parametrize("extra", [[], ["--unique_ids"], ["--native_loader"]], ids=["direct", "unique_ids", "native_loader"])
def test_lr2_app_on_the_nvlink_backend(native_lib, tmp_path, extra):
    rc, out = launch(tmp_path, ["lr2", "--", "--backend", "nvlink", "--train", "synthetic", "--test", "synthetic",
                                "--features", "3000", "--num_epochs", "6", "--batch_size", "100", "--learning_rate", "0.5",
                                "--synthetic_rows", "400", "--trace_step_interval", "8"] + extra)
    assert rc == 0, out[-3000:]
    assert "Finish train." in out and re.search(r"Finish evaluate, auc: \(0\.\d+, 0\.\d+\)", out), out[-3000:]
    assert re.search(r"\[worker:0\] epoch: \d+, local step: \d+, global step: \d+, loss: [\d.]+", out), out[-3000:]
