"""Queues, runners, coordinator, batcher (R10, P20, P22): the host-side input pipeline.

The reference builds its pipelines from ``tf.FIFOQueue`` + ``enqueue_many``/``dequeue`` +
``tf.train.batch`` + ``QueueRunner`` threads + ``tf.train.Coordinator`` with
``RunOptions(timeout_in_ms=...)`` on the consumer (lr2.py:159-175,205-208,471-475;
input_pipeline_large_dataset.py:18-66; input_pipeline.py:60-91).  The pattern kept here:
**producer thread(s) -> bounded queue -> batcher -> consumer with timeout + cooperative stop.**
"""
from __future__ import annotations

import threading
import time
from collections import deque
from typing import Callable, Iterable, List, Optional, Sequence


class QueueClosedError(Exception):
    """Dequeue on a closed, drained queue / enqueue on a closed queue (tf.errors.OutOfRangeError / CancelledError)."""


class DeadlineExceededError(Exception):
    """A blocking queue op ran past its timeout (tf.errors.DeadlineExceededError)."""


class FIFOQueue:
    def __init__(self, capacity: int, name: str = "fifo_queue"):
        self.capacity, self.name = int(capacity), name
        self._q = deque()
        self._lock = threading.Lock()
        self._not_empty = threading.Condition(self._lock)
        self._not_full = threading.Condition(self._lock)
        self._closed = False
        self._cancel = False

    def size(self) -> int:
        with self._lock:
            return len(self._q)

    def is_closed(self) -> bool:
        return self._closed

    def enqueue(self, item, timeout: Optional[float] = None) -> None:
        self.enqueue_many([item], timeout)

    def enqueue_many(self, items: Sequence, timeout: Optional[float] = None) -> None:
        """Blocks while the queue is full; elements go in one at a time (like TF, a batch larger
        than the free space is admitted progressively)."""
        deadline = None if timeout is None else time.monotonic() + timeout
        with self._not_full:
            for it in items:
                while len(self._q) >= self.capacity and not self._closed:
                    remaining = None if deadline is None else deadline - time.monotonic()
                    if remaining is not None and remaining <= 0:
                        raise DeadlineExceededError(f"{self.name}: enqueue timed out")
                    self._not_full.wait(remaining)
                if self._closed:
                    raise QueueClosedError(f"{self.name} is closed")
                self._q.append(it)
                self._not_empty.notify()

    def dequeue(self, timeout: Optional[float] = None):
        return self.dequeue_many(1, timeout)[0]

    def dequeue_many(self, n: int, timeout: Optional[float] = None) -> List:
        deadline = None if timeout is None else time.monotonic() + timeout
        out = []
        with self._not_empty:
            while len(out) < n:
                while not self._q:
                    if self._closed:
                        if out:
                            # put the partial batch back: tf.train.batch never emits short batches
                            self._q.extendleft(reversed(out))
                        raise QueueClosedError(f"{self.name} is closed and has insufficient elements")
                    remaining = None if deadline is None else deadline - time.monotonic()
                    if remaining is not None and remaining <= 0:
                        self._q.extendleft(reversed(out))
                        raise DeadlineExceededError(f"{self.name}: dequeue timed out")
                    self._not_empty.wait(remaining)
                out.append(self._q.popleft())
                self._not_full.notify()
        return out

    def close(self, cancel_pending_enqueues: bool = False) -> None:
        with self._lock:
            self._closed = True
            if cancel_pending_enqueues:
                self._cancel = True
            self._not_empty.notify_all()
            self._not_full.notify_all()


class Coordinator:
    """``tf.train.Coordinator``: cooperative stop for pipeline threads."""

    def __init__(self):
        self._stop = threading.Event()
        self._exc = None
        self._threads: List[threading.Thread] = []

    def should_stop(self) -> bool:
        return self._stop.is_set()

    def request_stop(self, ex: Optional[BaseException] = None) -> None:
        if ex is not None and self._exc is None:
            self._exc = ex
        self._stop.set()

    def register_thread(self, t: threading.Thread) -> None:
        self._threads.append(t)

    def wait_for_stop(self, timeout: Optional[float] = None) -> bool:
        return self._stop.wait(timeout)

    def join(self, threads: Optional[Iterable[threading.Thread]] = None, stop_grace_period_secs: float = 10.0):
        for t in list(threads if threads is not None else self._threads):
            t.join(stop_grace_period_secs)
            if t.is_alive():
                raise RuntimeError(f"thread {t.name} did not stop within {stop_grace_period_secs}s")
        if self._exc is not None:
            raise self._exc


class QueueRunner:
    """Runs ``enqueue_fn()`` in a loop on N threads until the coordinator stops or the queue closes."""

    def __init__(self, queue: FIFOQueue, enqueue_fns: Sequence[Callable[[], None]]):
        self.queue, self.enqueue_fns = queue, list(enqueue_fns)

    def create_threads(self, coord: Coordinator, daemon: bool = True, start: bool = True):
        threads = []
        for i, fn in enumerate(self.enqueue_fns):
            t = threading.Thread(target=self._run, args=(fn, coord), daemon=daemon,
                                 name=f"{self.queue.name}-runner-{i}")
            coord.register_thread(t)
            threads.append(t)
            if start:
                t.start()
        return threads

    def _run(self, fn, coord: Coordinator):
        try:
            while not coord.should_stop():
                fn()
        except (QueueClosedError, StopIteration):
            pass
        except BaseException as e:  # surface producer failures at coord.join()
            coord.request_stop(e)


_RUNNERS: List[QueueRunner] = []


def add_queue_runner(qr: QueueRunner) -> None:
    _RUNNERS.append(qr)


def start_queue_runners(coord: Coordinator, runners: Optional[Sequence[QueueRunner]] = None):
    """``tf.train.start_queue_runners``: start every registered runner; returns the threads."""
    threads = []
    for qr in (runners if runners is not None else list(_RUNNERS)):
        threads += qr.create_threads(coord)
    if runners is None:
        _RUNNERS.clear()
    return threads


class Batcher:
    """``tf.train.batch([dequeue_op], batch_size, capacity)``: a runner thread moves elements from
    ``source`` into an internal queue; ``next(timeout)`` returns exactly ``batch_size`` of them."""

    def __init__(self, source: FIFOQueue, batch_size: int, capacity: Optional[int] = None, name: str = "batch"):
        self.source, self.batch_size = source, batch_size
        self.queue = FIFOQueue(capacity or batch_size, name=name)
        self._held = None
        self.runner = QueueRunner(self.queue, [self._pump_safe])
        add_queue_runner(self.runner)

    def _pump_safe(self):
        """One bounded-wait move source -> batch queue; never blocks longer than 0.2 s so a
        coordinator stop is honoured promptly."""
        if self._held is None:
            try:
                self._held = (self.source.dequeue(timeout=0.2),)
            except DeadlineExceededError:
                return
            except QueueClosedError:
                self.queue.close()          # propagate end-of-input to the consumer
                raise
        try:
            self.queue.enqueue(self._held[0], timeout=0.2)
            self._held = None
        except DeadlineExceededError:
            pass

    def next(self, timeout: Optional[float] = None) -> List:
        return self.queue.dequeue_many(self.batch_size, timeout)


def batch(source: FIFOQueue, batch_size: int, capacity: Optional[int] = None, name: str = "batch") -> Batcher:
    return Batcher(source, batch_size, capacity, name)


def slice_input_producer(columns: Sequence[Sequence], shuffle: bool = False, capacity: int = 32,
                         num_epochs: Optional[int] = None, seed: Optional[int] = None,
                         name: str = "input_producer") -> FIFOQueue:
    """``tf.train.slice_input_producer``: a queue fed with one row ``(col0[i], col1[i], ...)`` at a
    time, cycling over the data (``num_epochs=None`` = forever), optionally reshuffled per epoch."""
    import random
    n = len(columns[0])
    q = FIFOQueue(capacity, name=name)
    state = {"epoch": 0, "order": list(range(n)), "pos": 0}
    rng = random.Random(seed)

    def produce():
        if state["pos"] == 0 and shuffle:
            rng.shuffle(state["order"])
        i = state["order"][state["pos"]]
        q.enqueue(tuple(c[i] for c in columns), timeout=0.2) if not q.is_closed() else None
        state["pos"] += 1
        if state["pos"] == n:
            state["pos"] = 0
            state["epoch"] += 1
            if num_epochs is not None and state["epoch"] >= num_epochs:
                q.close()
                raise StopIteration

    def produce_safe():
        try:
            produce()
        except DeadlineExceededError:
            pass

    add_queue_runner(QueueRunner(q, [produce_safe]))
    return q
