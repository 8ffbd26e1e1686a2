"""LR scratch / debug driver — the capability of the reference's lr2_debug.py.

The live part of that script (lr2_debug.py:1-231) is a bring-up tool for the data layer of the LR
trainer: the flag set with ``--train_file_list/--test_file_list`` (a *file of file names*, built by
run.sh), a loader that only counts lines, a ``multiprocessing`` loader wrapper the author could not
use ("gfile cannot run under multiprocessing", lr2_debug.py:115), ``GetFileList`` with worker
sharding ``lines[task_index::num_workers]`` (lr2_debug.py:220-224), and cluster/server creation
(lr2_debug.py:226-231).  Its dead part holds earlier trainers whose distinguishing ideas —
``init_weights(stddev=0.01)``, ``reduce_sum`` loss, CPU pinning, SyncReplicasOptimizerV2 and the
commented wide & deep extension (lr2_debug.py:288-289,404,423-435) — are real options here:

    python -m b200ps.apps.lr2_debug --train_file_list data/train_file_list --test_file_list data/test_file_list
        [--loader thread|process]           count lines of this worker's shard of the listed files
        [--train_steps N --deep_dim 128]    then run N steps of the (wide & deep) trainer on them (GPU)

This is a single-role tool: it loads ``cluster_conf.json`` for the worker count and the sharding, it
does not start ps tasks (use ``b200ps.apps.lr2`` for cluster runs).
"""
from __future__ import annotations

import multiprocessing
import os
import threading
import time
from typing import List

from .. import flags
from ..cluster.spec import ClusterSpec
from ..data.libsvm import GFile, Sample, count_lines
from ..io.log import info

FLAGS = flags.FLAGS
flags.define_role_flags(job_name_default="worker")
flags.DEFINE_integer("thread_num", 2, "thread num")
flags.DEFINE_float("learning_rate", 0.001, "Initial learning rate.")
flags.DEFINE_integer("num_epochs", 120, "Number of epochs to run trainer.")
flags.DEFINE_integer("batch_size", 500, "Batch size.")
flags.DEFINE_integer("features", 4762348, "Feature size")
flags.DEFINE_string("train_file_list", "data/train_file_list", "file holding one train path per line (lr2_debug.py:30)")
flags.DEFINE_string("test_file_list", "data/test_file_list", "file holding one test path per line (lr2_debug.py:31)")
flags.DEFINE_string("loader", "thread", "thread (lr2_debug.py:76-102) or process (lr2_debug.py:53-73)")
flags.DEFINE_integer("train_steps", 0, "after counting, run this many trainer steps on the listed train files (GPU)")
flags.DEFINE_integer("deep_dim", 128, "embedding width of the deep part (0 = wide only, i.e. plain LR)")
flags.DEFINE_float("init_stddev", 0.01, "init_weights stddev (lr2_debug.py:288-289)")
flags.DEFINE_boolean("loss_sum", False, "reduce_sum instead of reduce_mean loss (lr2_debug.py:432): lr is applied "
                     "to the summed gradient, i.e. multiplied by the batch size")
flags.DEFINE_integer("pin_cpu", -1, "pin this process to one CPU (the dead trainer pins ops to /cpu:0, lr2_debug.py:404)")


class LoadDataThread(threading.Thread):
    """Counts the lines of the files dealt to this thread (lr2_debug.py:76-102)."""

    def __init__(self, tid: int, files: List[str]):
        super().__init__()
        self.tid, self.files, self.line_num = tid, files, 0

    def run(self):
        for f in self.files:
            n = count_lines(f)
            info("thread: %d, file: %s, lines: %d" % (self.tid, f, n))
            self.line_num += n


def _count_in_process(files: List[str], out) -> None:
    out.put(sum(count_lines(f) for f in files))


class LoadDataProcess:
    """The multiprocessing wrapper of lr2_debug.py:53-73 — usable here because the file layer is plain
    Python I/O (the reference's gfile/JNI HDFS client did not survive a fork)."""

    def __init__(self, pid: int, files: List[str]):
        self.pid_, self.files, self.line_num = pid, files, 0
        self._q = multiprocessing.get_context("spawn").Queue()
        self._p = multiprocessing.get_context("spawn").Process(target=_count_in_process, args=(files, self._q))

    def start(self):
        self._p.start()

    def join(self):
        self.line_num = self._q.get()
        self._p.join()


class DataProvider:
    """Skeleton provider of lr2_debug.py:105-224: resolves the file lists and deals files to loaders."""

    def __init__(self, num_workers: int, task_index: int, thread_num: int, loader: str = "thread"):
        self.num_workers, self.task_index, self.thread_num, self.loader = num_workers, task_index, thread_num, loader

    def GetFileList(self, list_path: str) -> List[str]:
        with GFile(list_path) as f:
            lines = [line.strip() for line in f if line.strip()]
        return lines[self.task_index::self.num_workers]

    def CountLines(self, files: List[str]) -> int:
        cls = LoadDataProcess if self.loader == "process" else LoadDataThread
        workers = [cls(t, files[t::self.thread_num]) for t in range(self.thread_num) if files[t::self.thread_num]]
        for w in workers:
            w.start()
        for w in workers:
            w.join()
        return sum(w.line_num for w in workers)


def _train(files: List[str], test_files: List[str]) -> None:
    import torch
    from ..metrics.auc import StreamingAUC
    from ..parallel.context import init_from_env
    from ..train import SparseLRTrainer, SparseWideDeepTrainer
    samples = []
    for path in files:
        with GFile(path) as f:
            for line in f:
                if line.strip():
                    samples.append(Sample().parse_line_libsvm(line))
    if not samples:
        info("no training samples")
        return
    ctx = init_from_env()
    B = FLAGS.batch_size
    lr = FLAGS.learning_rate * (B if FLAGS.loss_sum else 1.0)
    if FLAGS.deep_dim > 0:
        tr = SparseWideDeepTrainer(ctx, FLAGS.features, dim=FLAGS.deep_dim, lr=lr, mode="sync" if FLAGS.sync else "async",
                                   use_graph=False)
    else:
        from ..models.sparse_lr import SparseLR
        tr = SparseLRTrainer(ctx, FLAGS.features, lr=lr, mode="sync" if FLAGS.sync else "async", use_graph=False,
                             model=SparseLR(FLAGS.features, stddev=FLAGS.init_stddev))

    def batch_of(chunk):
        return tuple(torch.from_numpy(a).to(tr.device) for a in Sample.format_samples_csr(chunk))

    t0 = time.time()
    for step in range(FLAGS.train_steps):
        lo = (step * B) % max(1, len(samples) - B + 1)
        tr.step(batch_of(samples[lo:lo + B]))
        if (step + 1) % max(1, FLAGS.train_steps // 5) == 0:
            info("local step: %d, global step: %d, loss: %.6f" % (step + 1, tr.global_step, tr.read_loss()))
    info("train time: %.3fs" % (time.time() - t0))
    test = []
    for path in test_files:
        with GFile(path) as f:
            for line in f:
                if line.strip():
                    test.append(Sample().parse_line_libsvm(line))
    if test:
        auc = StreamingAUC(200, device=tr.device)
        for lo in range(0, len(test), B):
            b = batch_of(test[lo:lo + B])
            pred, _ = tr.predict(b)
            auc.update(pred, b[3])
        info("Finish evaluate, auc: {}".format(auc.result()))


def main(argv=None) -> int:
    if FLAGS.pin_cpu >= 0 and hasattr(os, "sched_setaffinity"):
        os.sched_setaffinity(0, {FLAGS.pin_cpu})
    cluster_spec = ClusterSpec.from_json(FLAGS.cluster_conf)
    num_workers = cluster_spec.num_workers
    info("cluster: %s, num_workers: %d" % (cluster_spec.as_dict(), num_workers))
    dp = DataProvider(num_workers, FLAGS.task_index, FLAGS.thread_num, FLAGS.loader)
    train_files = dp.GetFileList(FLAGS.train_file_list)
    test_files = dp.GetFileList(FLAGS.test_file_list) if FLAGS.test_file_list and os.path.exists(FLAGS.test_file_list) else []
    info("train files: %s" % ",".join(train_files))
    info("test files: %s" % ",".join(test_files))
    info("train lines: %d" % dp.CountLines(train_files))
    info("test lines: %d" % dp.CountLines(test_files))
    if FLAGS.train_steps > 0:
        _train(train_files, test_files)
    return 0


if __name__ == "__main__":
    flags.run(main)
