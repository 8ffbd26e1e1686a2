"""Feeding-thread queue pipeline demo — the reference's input_pipeline_large_dataset.py.

A python feeder thread pushes wrap-around chunks of 20 rows of a 100 003 x 4 synthetic table
(plus one-hot targets) into a FIFO of capacity 50; a batcher assembles batches of 15
(capacity 40); the consumer dequeues with a 4 s timeout (``RunOptions(timeout_in_ms=4000)``)
and, unlike the reference (which calls ``exit(0)`` before its shutdown code,
input_pipeline_large_dataset.py:68-74), stops the pipeline cleanly.
"""
from __future__ import annotations

import threading

import numpy as np

from .. import flags
from ..data.queue import Batcher, Coordinator, DeadlineExceededError, FIFOQueue, QueueClosedError, start_queue_runners

FLAGS = flags.FLAGS
flags.DEFINE_integer("pipeline_steps", 1000, "number of batches to dequeue")
flags.DEFINE_integer("pipeline_print_every", 100, "print every Nth batch")


def main(argv=None) -> int:
    r = np.arange(0.0, 100003.0)
    raw_data = np.dstack((r, r, r, r))[0]
    raw_target = np.array([[1, 0, 0]] * 100003)
    queue = FIFOQueue(capacity=50, name="input_queue")
    batcher = Batcher(queue, batch_size=15, capacity=40, name="batch15")
    coord = Coordinator()
    stop = threading.Event()

    def enqueue():
        """Iterates over the data in chunks of 20, wrapping around at the end (…:29-50)."""
        under = 0
        mx = len(raw_data)
        while not stop.is_set():
            upper = under + 20
            if upper <= mx:
                curr_data, curr_target = raw_data[under:upper], raw_target[under:upper]
                under = upper
            else:
                rest = upper - mx
                curr_data = np.concatenate((raw_data[under:mx], raw_data[0:rest]))
                curr_target = np.concatenate((raw_target[under:mx], raw_target[0:rest]))
                under = rest
            try:
                queue.enqueue_many(list(zip(curr_data, curr_target)), timeout=0.5)
            except DeadlineExceededError:
                continue
            except QueueClosedError:
                break
        print("finished enqueueing")

    t = threading.Thread(target=enqueue, daemon=True)
    t.start()
    threads = start_queue_runners(coord)
    for i in range(FLAGS.pipeline_steps):
        rows = batcher.next(timeout=4.0)
        if i % FLAGS.pipeline_print_every == 0:
            data = np.stack([d for d, _ in rows])
            print(i, data[:, 0].tolist())
    stop.set()
    queue.close(cancel_pending_enqueues=True)
    coord.request_stop()
    coord.join(threads)
    t.join(timeout=5)
    print("pipeline stopped cleanly")
    return 0


if __name__ == "__main__":
    flags.run(main)
