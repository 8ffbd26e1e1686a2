"""Sparse logistic-regression parameter-server trainer — the capability of the reference's lr2.py.

    python -m b200ps.apps.lr2 --job_name=ps --task_index=0
    python -m b200ps.apps.lr2 --job_name=worker --task_index=0 --train=a,b --test=c --mode=all

Every reference flag keeps its name and default (lr2.py:20-35); the cluster comes from
``cluster_conf.json`` in the working directory (lr2.py:325).  Flow per worker (lr2.py:348-476):
load data (``all``) or start loader threads (``queue``), build the model on the parameter
servers, wait for the chief's initialisation, train ``num_epochs`` epochs with a test-loss
trace every ``trace_step_interval`` steps and at each epoch end, then a streaming-AUC pass over
the test data and queue teardown.

``--backend nvlink``: W is row-sharded over the GPUs and each step is the fused
gather/loss/scatter kernel (K4/K5); ``--backend gloo``: ids go up to the ps, rows come down,
``IndexedSlices`` go back up — the reference's own message pattern, on CPU.
``--train=synthetic`` / ``--test=synthetic`` generate data instead of reading files.
"""
from __future__ import annotations

import os
import tempfile

import torch

from .. import flags
from ..cluster.server import Server
from ..cluster.spec import ClusterSpec
from ..cluster.supervisor import Supervisor
from ..data.libsvm import Sample, csr_from_sparse_triple
from ..data.provider import DataProvider
from ..data.queue import DeadlineExceededError, QueueClosedError
from ..io.log import debug, error, info  # noqa: F401  (the reference's helpers, lr2.py:38-48)
from ..metrics.auc import StreamingAUC
from ..models.sparse_lr import SparseLR, sparse_lr_reference
from ..models.wide_deep import SparseWideDeep, wide_deep_reference

FLAGS = flags.FLAGS
flags.define_role_flags(job_name_default="worker")
flags.DEFINE_integer("thread_num", 2, "thread num")
flags.DEFINE_float("learning_rate", 0.001, "Initial learning rate.")
flags.DEFINE_integer("num_epochs", 120, "Number of epochs to run trainer.")
flags.DEFINE_integer("batch_size", 500, "Batch size. Must divide evenly into the dataset sizes.")
flags.DEFINE_integer("features", 4762348, "Feature size")
flags.DEFINE_string("train", "hdfs://localhost:9000/user/yaowq/tensorflow/lr/data/train/part-00000", "train file", short="t")
flags.DEFINE_string("test", "hdfs://localhost:9000/user/yaowq/tensorflow/lr/data/test/part-00000", "test file", short="T")
flags.DEFINE_string("checkpoint", "hdfs://localhost:9000/user/yaowq/tensorflow/lr/checkpoint",
                    "checkpoint file (declared and unused in the reference; use --logdir to checkpoint)")
flags.DEFINE_integer("trace_step_interval", 10000, "number of steps to output info")
flags.DEFINE_float("train_sampling_rate", 1.0, "samling rate for train file")
flags.DEFINE_float("test_sampling_rate", 1.0, "samling rate for test file")
flags.DEFINE_string("mode", "all", "load data all or queue")
flags.DEFINE_integer("train_queue_capacity", 2000, "train queue capacity")
flags.DEFINE_integer("test_queue_capacity", 2000, "test queue capacity")
flags.DEFINE_string("train_file_list", "", "file holding one train path per line (lr2_debug.py:30)")
flags.DEFINE_string("test_file_list", "", "file holding one test path per line (lr2_debug.py:31)")
flags.DEFINE_integer("deep_dim", 0, "0 = plain LR (lr2.py); 128/256/512 adds the [features,dim] embedding + [dim,1] "
                     "deep part sketched at lr2_debug.py:423-428")
flags.DEFINE_float("init_stddev", 0.0, "stddev of the weight initialiser; 0 = the script's default (1.0 for LR as in "
                   "lr2.py:384, 0.01 with --deep_dim as in lr2_debug.py:288-289)")
flags.DEFINE_boolean("native_loader", False, "read, sample and batch the train files with the native loader threads instead "
                     "of the Python DataProvider (GPU backend: packed pinned batches -> one H2D copy + one graph launch "
                     "per step; gloo backend: the CSR arrays go straight to the CPU step)")
flags.DEFINE_boolean("unique_ids", False, "GPU backend, plain LR: make each batch's ids unique on the device first (one read and "
                     "one update per distinct id, as embedding_lookup_sparse / IndexedSlices do); pays when ids repeat")
flags.DEFINE_integer("synthetic_rows", 4000, "rows per synthetic file when --train/--test is 'synthetic'")


def _csr(samples):
    """A batch as CPU tensors (row_ptr, fid, fval, labels): a list of Samples is formatted, a CSR tuple (numpy or
    torch: what the native loader hands out) is taken as it is."""
    if isinstance(samples, tuple):
        return tuple(a if isinstance(a, torch.Tensor) else torch.from_numpy(a) for a in samples)
    return tuple(torch.from_numpy(a) for a in Sample.format_samples_csr(samples))


def _to_device_batch(samples, device):
    return tuple(a.to(device) for a in _csr(samples))


class _CpuRunner:
    """gloo path: unique ids up, rows down, gradient slices up — lr2.py:390,396 as messages."""

    def __init__(self, store, model, lr, sync, replicas):
        self.store, self.model, self.lr, self.sync, self.replicas = store, model, lr, sync, replicas

    def _fwd(self, samples, train):
        row_ptr, fid, fval, labels = _csr(samples)
        uniq, inv = torch.unique(fid, return_inverse=True)
        rows = self.store.sparse_pull(self.model.W, uniq)
        self.store.pull()
        b = self.store.var(self.model.Bn)
        loss, pred, dW, db = sparse_lr_reference(rows, b, row_ptr, inv, fval, labels, train=train)
        return loss, pred, dW, db, uniq, labels

    def train_step(self, samples):
        loss, pred, dW, db, uniq, _ = self._fwd(samples, True)
        g = torch.zeros(uniq.numel(), dtype=torch.float64).index_add_(0, dW[0], dW[1])
        scale = 1.0 / self.replicas if self.sync else 1.0
        self.store.sparse_push(self.model.W, uniq, g.float(), self.lr * scale)
        self.store.grad(self.model.Bn).fill_(float(db))
        if self.sync:
            self.store.push_sync(self.lr, self.replicas)
        else:
            self.store.push_async(self.lr)
        return float(loss)

    def eval_batch(self, samples):
        loss, pred, _, _, _, labels = self._fwd(samples, False)
        return float(loss), pred.float(), labels


class _CpuWideDeepRunner:
    """gloo path of the wide & deep extension: unique ids up, W/E rows down, w2 whole, four pushes."""

    def __init__(self, store, model, lr, sync, replicas):
        self.store, self.model, self.lr, self.sync, self.replicas = store, model, lr, sync, replicas
        self._all = torch.arange(model.dim)

    def _fwd(self, samples, train):
        m = self.model
        row_ptr, fid, fval, labels = _csr(samples)
        uniq, inv = torch.unique(fid, return_inverse=True)
        W = self.store.sparse_pull(m.W, uniq) if m.wide else None
        E = self.store.sparse_pull(m.E, uniq)
        w2 = self.store.sparse_pull(m.W2, self._all)
        self.store.pull()
        b = self.store.var(m.Bn)
        loss, pred, grads = wide_deep_reference(W, E, w2, b, row_ptr, inv, fval, labels, train=train)
        return loss, pred, grads, uniq, labels

    def train_step(self, samples):
        m = self.model
        loss, pred, g, uniq, _ = self._fwd(samples, True)
        lr = self.lr * (1.0 / self.replicas if self.sync else 1.0)
        if m.wide:
            self.store.sparse_push(m.W, uniq, g["W"].float(), lr)
        self.store.sparse_push(m.E, uniq, g["E"].float(), lr)
        self.store.sparse_push(m.W2, self._all, g["w2"].float(), lr)
        self.store.grad(m.Bn).fill_(float(g["b"]))
        if self.sync:
            self.store.push_sync(self.lr, self.replicas)
        else:
            self.store.push_async(self.lr)
        return float(loss)

    def eval_batch(self, samples):
        loss, pred, _, _, labels = self._fwd(samples, False)
        return float(loss), pred.float(), labels


class _GpuRunner:
    def __init__(self, trainer):
        self.tr = trainer

    def train_step(self, samples):
        hb = FLAGS.heartbeat_steps
        if hb > 0 and self.tr.local_step % hb == 0:
            self.tr.ctx.check_peers(min(FLAGS.barrier_timeout_secs, 10.0))      # raises PeerLost (restart + restore)
        self.tr.step(_to_device_batch(samples, self.tr.device))
        return None

    def eval_batch(self, samples):
        batch = _to_device_batch(samples, self.tr.device)
        pred, loss = self.tr.predict(batch)
        return float(loss.item()), pred, batch[3]


def Test(runner, store, iter_num, step_num, test_samples):
    """lr2.py:307-315 — loss on the cached test sample + the shared global step."""
    if not test_samples:
        return None
    loss, _, _ = runner.eval_batch(test_samples)
    info("epoch: {}, local step: {}, global step: {}, loss: {}".format(iter_num, step_num, store.global_step, loss))
    return loss


def _maybe_synthetic(spec: str, kind: str, task_index: int) -> str:
    if spec != "synthetic":
        return spec
    from ..data.synthetic import write_libsvm_files
    d = os.path.join(tempfile.gettempdir(), f"b200ps_synth_{os.getuid()}_{kind}")
    paths = write_libsvm_files(d, 2 if kind == "train" else 1, FLAGS.synthetic_rows, nnz_mean=20, spread=5,
                               features=min(FLAGS.features, 1 << 20), seed=5 + (kind == "test"))
    return ",".join(paths)


def _file_list(path: str) -> str:
    from ..data.libsvm import GFile
    with GFile(path) as f:
        return ",".join(line.strip() for line in f if line.strip())


def main(argv=None) -> int:
    learning_rate, num_epochs = FLAGS.learning_rate, FLAGS.num_epochs
    trace_step_interval = FLAGS.trace_step_interval
    cluster_spec = ClusterSpec.from_json(FLAGS.cluster_conf)
    num_workers = cluster_spec.num_workers
    server = Server(cluster_spec, job_name=FLAGS.job_name, task_index=FLAGS.task_index, backend=FLAGS.backend, timeout_s=FLAGS.barrier_timeout_secs)

    if FLAGS.job_name == "ps":
        info("start ...")
        server.join()
        return 0
    if FLAGS.job_name != "worker":
        return 2

    info("start ...")
    is_chief = FLAGS.task_index == 0
    if FLAGS.train_file_list:
        FLAGS.train = _file_list(FLAGS.train_file_list)
    if FLAGS.test_file_list:
        FLAGS.test = _file_list(FLAGS.test_file_list)
    FLAGS.train = _maybe_synthetic(FLAGS.train, "train", FLAGS.task_index)
    FLAGS.test = _maybe_synthetic(FLAGS.test, "test", FLAGS.task_index)
    data_provider = DataProvider.from_flags(FLAGS, num_workers)
    if FLAGS.mode == "all":
        info("load data")
        data_provider.init()
        data_provider.LoadData()

    info("build graph")
    gpu = server.backend == "nvlink"
    deep = FLAGS.deep_dim > 0
    if deep:
        model = SparseWideDeep(FLAGS.features, dim=FLAGS.deep_dim, stddev=FLAGS.init_stddev or 0.01)
    else:
        model = SparseLR(FLAGS.features, stddev=FLAGS.init_stddev or 1.0, dedupe=True if FLAGS.unique_ids else None)
    sync = bool(FLAGS.sync)
    if gpu:
        from ..train import SparseLRTrainer
        ctx = server.nvlink_context()
        trainer = SparseLRTrainer(ctx, FLAGS.features, lr=learning_rate, mode="sync" if sync else "async",
                                  placement=FLAGS.placement, seed=FLAGS.seed, use_graph=False, model=model)
        store, runner = trainer.store, _GpuRunner(trainer)
    else:
        from ..ps.gloo_store import GlooParameterStore
        ctx = None
        store = GlooParameterStore(server, placement=FLAGS.placement, seed=FLAGS.seed)
        model.register(store)
        store.build()
        runner = (_CpuWideDeepRunner if deep else _CpuRunner)(store, model, learning_rate, sync,
                                                              FLAGS.replicas_to_aggregate or num_workers)
    supervisor = Supervisor(is_chief=is_chief, store=store, ctx=ctx, logdir=FLAGS.logdir or None,
                            save_model_secs=FLAGS.save_secs, ready_timeout_secs=FLAGS.barrier_timeout_secs)

    info("Start session ...")
    with supervisor.prepare_or_wait_for_session(server.target):
        # (the reference re-runs init_op here on EVERY worker, lr2.py:419, resetting the model
        #  whenever a worker joins late; only the chief initialises in this framework)
        if FLAGS.mode == "queue":
            info("load data queue ...")
            data_provider.init()
            data_provider.LoadData()

        info("sampling test data ...")
        _, _, _, _, n_test = t = data_provider.GetTestSamplesSampled(sampling_rate=0.1, sampling_max_num=1000)
        # keep the sampled rows as Sample objects for the periodic trace
        test_samples = []
        if n_test:
            labels, fids, fvals, sp_indices, _ = t
            rp, fid, fv, y = csr_from_sparse_triple(sp_indices, fids, fvals, labels)
            for i in range(n_test):
                s = Sample()
                s.label = int(y[i]); s.indices = fid[rp[i]:rp[i + 1]].tolist(); s.values = fv[rp[i]:rp[i + 1]].tolist()
                test_samples.append(s)

        info("Start train ...")
        step_num, iter_num = 0, 0
        info("num_epochs: %d" % num_epochs)
        if FLAGS.native_loader:
            # the input side of lr2.py:87-169 in native threads: thread_num readers over this worker's files,
            # Bernoulli sampling, batches of exactly batch_size rows (pinned buffers on the GPU backend)
            from ..data.native_loader import NativeLibsvmLoader
            files = DataProvider.GetFileList(FLAGS.train, num_workers, FLAGS.task_index)
            while iter_num < num_epochs:
                loader = NativeLibsvmLoader(files, FLAGS.batch_size, threads=FLAGS.thread_num,
                                            sampling_rate=FLAGS.train_sampling_rate, epochs=1, seed=FLAGS.seed + iter_num,
                                            pinned=gpu)
                base = step_num

                def on_loss(i, loss, base=base, epoch=iter_num):
                    if (base + i + 1) % trace_step_interval == 0:
                        Test(runner, store, epoch, base + i + 1, test_samples)
                if gpu:
                    step_num += trainer.fit_from_loader(loader, on_loss=on_loss)
                else:
                    for i, batch in enumerate(loader):
                        n, z = batch.rows, batch.nnz
                        rp, fid, fval, y = batch.arrays()
                        runner.train_step((rp[: n + 1].clone(), fid[:z].clone(), fval[:z].clone(), y[:n].clone()))
                        batch.release()
                        step_num += 1
                        on_loss(i, None)
                info("native loader: %s" % loader.stats())
                loader.close()
                Test(runner, store, iter_num, step_num, test_samples)
                iter_num += 1
        while iter_num < num_epochs:
            data_provider.Shuffle()
            try:
                for batch_samples in data_provider.NextBatch(data_type="train"):
                    if batch_samples is None or len(batch_samples) <= 0:
                        break
                    if FLAGS.fault_inject_step >= 0 and step_num == FLAGS.fault_inject_step \
                            and FLAGS.task_index == FLAGS.fault_rank:
                        error("fault injection: worker exiting abruptly")
                        os._exit(17)
                    runner.train_step(batch_samples)
                    step_num += 1
                    if step_num % trace_step_interval == 0:
                        Test(runner, store, iter_num, step_num, test_samples)
            except (DeadlineExceededError, QueueClosedError) as e:
                print("ignore exception: ", e)
            Test(runner, store, iter_num, step_num, test_samples)
            iter_num += 1
        info("Finish train.")

        info("Start evaluate ...")
        auc = StreamingAUC(200, device=(trainer.device if gpu else "cpu"))
        auc_val = None
        try:
            for batch_samples in data_provider.NextBatch(data_type="test"):
                if batch_samples is None or len(batch_samples) <= 0:
                    break
                _, pred, labels = runner.eval_batch(batch_samples)
                auc.update(pred, labels)
                auc_val = (auc.result(), auc.result())     # (value, update_op) like the reference's tuple
        except (DeadlineExceededError, QueueClosedError) as e:
            print("ignore exception: ", e)
        info("Finish evaluate, auc: {}".format(auc_val))
        if is_chief and FLAGS.logdir:
            supervisor.save()

    if FLAGS.mode == "queue":
        data_provider.close()
    supervisor.stop()
    server.post_done()
    return 0


if __name__ == "__main__":
    flags.run(main)
