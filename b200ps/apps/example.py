"""MNIST sigmoid-MLP parameter-server trainer — the capability of the reference's example.py.

    pc-01$ python -m b200ps.apps.example --job_name=ps --task_index=0
    pc-02$ python -m b200ps.apps.example --job_name=worker --task_index=0
    pc-03$ python -m b200ps.apps.example --job_name=worker --task_index=1

Same roles, flags, default cluster (1 ps + 2 workers on localhost:2222-2224, example.py:24-26),
hyper-parameters (batch 100, lr 0.0005, 5 epochs, example.py:43-45), log directory
(``./logs/mnist/{job}_{task}``), summaries (``cost``, ``accuracy``) and console lines
(example.py:178-189,192).  ``--sync`` turns on the synchronous-replica variant the reference
leaves commented out (example.py:109-123,147-152).

``--backend gloo`` (default without a GPU): real ps process, CPU math, gloo/TCP transport.
``--backend nvlink``: every worker drives one B200 and is also a ps shard; all hot kernels
are native (libb200ps.so).  MNIST itself cannot be downloaded here: ``MNIST_data/{job}_{task}``
is used when it holds the idx files, otherwise a synthetic 28x28 problem of the same shape.
"""
from __future__ import annotations

import os
import shutil
import sys
import time

import torch

from .. import flags
from ..cluster.server import Server
from ..cluster.spec import ClusterSpec
from ..cluster.supervisor import Supervisor
from ..data.synthetic import read_data_sets
from ..io import log
from ..io.events import SummaryWriter
from ..models.mlp import SigmoidMLP, mlp_reference_step

FLAGS = flags.FLAGS
flags.define_role_flags(job_name_default="")
flags.DEFINE_string("ps_hosts", "localhost:2222", "comma-separated ps tasks (example.py:24)")
flags.DEFINE_string("worker_hosts", "localhost:2223,localhost:2224", "comma-separated worker tasks (example.py:25-26)")
flags.DEFINE_integer("batch_size", 100, "per-worker batch (example.py:43)")
flags.DEFINE_float("learning_rate", 0.0005, "SGD step (example.py:44)")
flags.DEFINE_integer("training_epochs", 5, "epochs (example.py:45; the docs say 20)")
flags.DEFINE_string("logs_root", "./logs/mnist", "summary root; the task dir is {job}_{task} (example.py:48)")
flags.DEFINE_string("data_root", "MNIST_data", "MNIST idx root; the task dir is {job}_{task} (example.py:61)")
flags.DEFINE_integer("train_examples", 55000, "synthetic training-set size (MNIST: 55000)")
flags.DEFINE_integer("test_examples", 10000, "synthetic test-set size (MNIST: 10000)")
flags.DEFINE_integer("frequency", 100, "progress line every N batches (example.py:144)")
flags.DEFINE_integer("summary_every", 1, "write summaries every N steps (reference: every step)")


def build_cluster() -> ClusterSpec:
    if FLAGS.cluster_conf and os.path.exists(FLAGS.cluster_conf) and "--cluster_conf" in " ".join(sys.argv):
        return ClusterSpec.from_json(FLAGS.cluster_conf)
    spec = {"worker": [h for h in FLAGS.worker_hosts.split(",") if h]}
    ps = [h for h in FLAGS.ps_hosts.split(",") if h]
    if ps:
        spec["ps"] = ps
    return ClusterSpec(spec)


def main(argv=None) -> int:
    cluster = build_cluster()
    if FLAGS.job_name not in ("ps", "worker"):
        print("--job_name must be 'ps' or 'worker'", file=sys.stderr)
        return 2
    server = Server(cluster, FLAGS.job_name, FLAGS.task_index, backend=FLAGS.backend, timeout_s=FLAGS.barrier_timeout_secs)
    batch_size, learning_rate, training_epochs = FLAGS.batch_size, FLAGS.learning_rate, FLAGS.training_epochs
    logs_path = os.path.join(FLAGS.logs_root, "{}_{}".format(FLAGS.job_name, FLAGS.task_index))
    if os.path.exists(logs_path):
        shutil.rmtree(logs_path)
    os.makedirs(logs_path)            # (the reference's bare os.mkdir fails when ./logs/mnist is missing)

    if FLAGS.job_name == "ps":
        print("ps {} start ...".format(FLAGS.task_index), flush=True)
        server.join()
        return 0

    print("worker {} start ...".format(FLAGS.task_index), flush=True)
    gpu = server.backend == "nvlink"
    device = f"cuda:{FLAGS.task_index % max(torch.cuda.device_count(), 1)}" if gpu else "cpu"
    data_path = os.path.join(FLAGS.data_root, "{}_{}".format(FLAGS.job_name, FLAGS.task_index))
    mnist = read_data_sets(data_path, one_hot=False, device=device, seed=FLAGS.seed + FLAGS.task_index)
    if mnist.train.num_examples != FLAGS.train_examples and hasattr(mnist.train, "seed"):
        mnist.train.num_examples, mnist.test.num_examples = FLAGS.train_examples, FLAGS.test_examples
    is_chief = FLAGS.task_index == 0
    mode = "sync" if FLAGS.sync else "async"
    R = FLAGS.replicas_to_aggregate or cluster.num_workers

    model = SigmoidMLP()
    if gpu:
        from ..train import MLPTrainer
        ctx = server.nvlink_context(device)
        trainer = MLPTrainer(ctx, batch_size=batch_size, lr=learning_rate, mode=mode, replicas_to_aggregate=R,
                             placement=FLAGS.placement, seed=FLAGS.seed, model=model)
        store = trainer.store
    else:
        from ..ps.gloo_store import GlooParameterStore
        ctx, trainer = None, None
        store = GlooParameterStore(server, placement=FLAGS.placement if FLAGS.placement != "flat_shard"
                                   else "round_robin_variable", seed=FLAGS.seed)
        model.register(store)
        store.build()
    print("Variables initialized ...", flush=True)

    sv = Supervisor(is_chief=is_chief, store=store, ctx=ctx, logdir=FLAGS.logdir or None,
                    save_model_secs=FLAGS.save_secs, ready_timeout_secs=FLAGS.barrier_timeout_secs)
    begin_time = time.time()
    frequency = FLAGS.frequency
    cost = float("nan")
    with sv.prepare_or_wait_for_session(server.target):
        log.info("go into sv.prepare")
        writer = SummaryWriter(logs_path)
        start_time = time.time()
        local_step = 0
        for epoch in range(training_epochs):
            batch_count = int(mnist.train.num_examples / batch_size)
            count = 0
            for i in range(batch_count):
                batch_x, batch_y = mnist.train.next_batch(batch_size)
                if FLAGS.fault_inject_step >= 0 and local_step == FLAGS.fault_inject_step \
                        and FLAGS.task_index == FLAGS.fault_rank:
                    log.error("fault injection: worker exiting abruptly")
                    os._exit(17)
                if gpu:
                    if FLAGS.heartbeat_steps > 0 and local_step % FLAGS.heartbeat_steps == 0:
                        ctx.check_peers(min(FLAGS.barrier_timeout_secs, 10.0))     # raises PeerLost (restart + restore)
                    trainer.step(batch_x, batch_y)
                    want_host = (local_step % FLAGS.summary_every == 0) or count + 1 == frequency or i + 1 == batch_count
                    if want_host:
                        cost = trainer.read_loss()
                        acc = trainer.last_correct.item() / batch_size
                        step = store.global_step
                else:
                    if mode == "async":
                        store.pull()
                    params = {n: store.var(n) for n in model.names()}
                    loss_t, grads, correct = mlp_reference_step(params, batch_x, batch_y, names=model.names())
                    for n in model.names():
                        store.grad(n).copy_(grads[n])
                    if mode == "sync":
                        store.push_sync(learning_rate, R)
                    else:
                        store.push_async(learning_rate)
                    cost, acc = float(loss_t), correct / batch_size
                    want_host = True
                    step = store.global_step if (local_step % FLAGS.summary_every == 0) else 0
                local_step += 1
                if want_host and local_step % FLAGS.summary_every == 0:
                    writer.add_summary({"cost": cost, "accuracy": acc}, step)
                count += 1
                if count % frequency == 0 or i + 1 == batch_count:
                    elapsed_time = time.time() - start_time
                    start_time = time.time()
                    global_step_val = store.global_step
                    print("Step: %d," % (step + 1),
                          " Global Step: %2d," % (global_step_val),
                          " Epoch: %2d," % (epoch + 1),
                          " Batch: %3d of %3d," % (i + 1, batch_count),
                          " Cost: %.4f," % cost,
                          " AvgTime: %3.2fms" % float(elapsed_time * 1000 / frequency), flush=True)
                    count = 0

        # full test set in one evaluation, like example.py:187
        if gpu:
            test_acc, _ = trainer.evaluate(mnist.test.images, mnist.test.labels)
        else:
            store.pull()
            params = {n: store.var(n) for n in model.names()}
            _, _, correct = mlp_reference_step(params, mnist.test.images, mnist.test.labels, names=model.names(), train=False)
            test_acc = correct / mnist.test.num_examples
        print("Test-Accuracy: %2.2f" % test_acc)
        print("Total Time: %3.2fs" % float(time.time() - begin_time))
        print("Final Cost: %.4f" % cost)
        writer.close()
        if is_chief and FLAGS.logdir:
            sv.save()

    sv.stop()
    server.post_done()
    print("done", flush=True)
    return 0


if __name__ == "__main__":
    flags.run(main)
