"""DataProvider (P19-P21): lr2.py:87-304 with the same public methods.

* file sharding by worker ``files[task_index::num_workers]`` (lr2.py:302-304) and by loader
  thread ``files[tid::thread_num]`` (lr2.py:232-233);
* ``mode='all'``: loader threads read their files fully, Bernoulli-subsample lines
  (``random() < 1 - rate`` drops, lr2.py:105) and the samples are merged after join;
  ``Shuffle()`` permutes the training list; ``NextBatch`` yields slices of ``batch_size``;
* ``mode='queue'``: daemon loader threads loop over their files forever and push raw lines
  into a bounded FIFO (capacity ``--train_queue_capacity``); a batcher assembles
  ``batch_size`` lines; ``NextBatch`` yields ONE batch per call (so an "epoch" is a step,
  exactly as in the reference).

Deliberate fixes of reference quirks: ``is_train`` is honoured, so test loader
threads feed the *test* queue and ``GetTestSamplesSampled`` reads the test queue
(lr2.py:233,226 mix them up); loading needs no session object.
"""
from __future__ import annotations

import random
import threading
import time
from typing import List, Optional

import numpy as np

from ..io import log
from .libsvm import GFile, Sample
from .queue import (Batcher, Coordinator, DeadlineExceededError, FIFOQueue, QueueClosedError,
                    start_queue_runners)


class LoadDataThread(threading.Thread):
    def __init__(self, tid: int, files: List[str], sampling_rate: float, mode: str = "all",
                 queue: Optional[FIFOQueue] = None, is_train: bool = True, batch_size: int = 500,
                 coord: Optional[Coordinator] = None, seed: Optional[int] = None):
        super().__init__(daemon=True, name=f"load-{'train' if is_train else 'test'}-{tid}")
        self.samples: List[Sample] = []
        self.tid, self.files, self.sampling_rate, self.mode = tid, files, sampling_rate, mode
        self.queue, self.is_train, self.batch_size, self.coord = queue, is_train, batch_size, coord
        self._rng = random.Random(seed)
        self.total_lines = 0

    def _keep(self, line: str) -> bool:
        return not (line is None or len(line) < 2 or self._rng.random() < 1 - self.sampling_rate)

    def run(self):
        if self.mode == "all":
            for file_name in self.files:
                log.info("thread: %d, input file: %s" % (self.tid, file_name))
                n = 0
                with GFile(file_name) as f:
                    for line in f:
                        if not self._keep(line):
                            continue
                        self.samples.append(Sample().parse_line_libsvm(line))
                        n += 1
                log.info("thread: %d, input file: %s, samples: %d" % (self.tid, file_name, n))
            log.info("thread: %d, samples: %d" % (self.tid, len(self.samples)))
            self.total_lines = len(self.samples)
        elif self.mode == "queue":
            lines: List[str] = []
            try:
                while self.files and not (self.coord and self.coord.should_stop()):
                    for file_name in self.files:
                        log.debug("thread: %d, input file: %s" % (self.tid, file_name))
                        with GFile(file_name) as f:
                            for line in f:
                                if not self._keep(line):
                                    continue
                                lines.append(line)
                                self.total_lines += 1
                                if len(lines) == self.batch_size:
                                    self._flush(lines)
                                    lines = []
                                if self.coord and self.coord.should_stop():
                                    return
            except QueueClosedError:
                return

    def _flush(self, lines: List[str]) -> None:
        while True:
            try:
                self.queue.enqueue_many(lines, timeout=0.5)
                return
            except DeadlineExceededError:
                if self.coord and self.coord.should_stop():
                    raise QueueClosedError("stopping")


class DataProvider:
    def __init__(self, num_workers: int, task_index: int, thread_num: int, mode: str = "all",
                 train: str = "", test: str = "", batch_size: int = 500, train_sampling_rate: float = 1.0,
                 test_sampling_rate: float = 1.0, train_queue_capacity: int = 2000,
                 test_queue_capacity: int = 2000, dequeue_timeout: float = 1.0, seed: Optional[int] = None):
        self.mode, self.num_workers, self.task_index, self.thread_num = mode, num_workers, task_index, thread_num
        self.train, self.test, self.batch_size = train, test, batch_size
        self.train_sampling_rate, self.test_sampling_rate = train_sampling_rate, test_sampling_rate
        self.dequeue_timeout = dequeue_timeout          # the RunOptions(timeout_in_ms=1000) of lr2.py:175
        self.seed = seed
        self.train_samples: List[Sample] = []
        self.test_samples: List[Sample] = []
        self.train_file_list: List[str] = []
        self.test_file_list: List[str] = []
        self.train_threads = None
        self.test_threads = None
        self.coord: Optional[Coordinator] = None
        self.threads_dequeue = None
        self.train_data_queue = FIFOQueue(train_queue_capacity, "train_data_queue")
        self.test_data_queue = FIFOQueue(test_queue_capacity, "test_data_queue")
        self.train_data_batch: Optional[Batcher] = None
        self.test_data_batch: Optional[Batcher] = None

    @classmethod
    def from_flags(cls, FLAGS, num_workers: int) -> "DataProvider":
        return cls(num_workers, FLAGS.task_index, FLAGS.thread_num, FLAGS.mode, FLAGS.train, FLAGS.test,
                   FLAGS.batch_size, FLAGS.train_sampling_rate, FLAGS.test_sampling_rate,
                   FLAGS.train_queue_capacity, FLAGS.test_queue_capacity)

    # -- lr2.py:194-208 -------------------------------------------------------------
    def init(self, sess=None) -> None:
        self.train_file_list = self.GetFileList(self.train, self.num_workers, self.task_index)
        self.test_file_list = self.GetFileList(self.test, self.num_workers, self.task_index)
        if self.mode == "queue":
            self.coord = Coordinator()
            self.train_data_batch = Batcher(self.train_data_queue, self.batch_size, self.batch_size, "train_batch")
            self.test_data_batch = Batcher(self.test_data_queue, self.batch_size, self.batch_size, "test_batch")
            self.threads_dequeue = start_queue_runners(self.coord)
        self.train_threads = self.InitThreads(self.train_file_list, len(self.train_file_list), self.thread_num,
                                              self.train_sampling_rate, is_train=True)
        self.test_threads = self.InitThreads(self.test_file_list, len(self.test_file_list), self.thread_num,
                                             self.test_sampling_rate, is_train=False)

    def GetTrainSamples(self):
        return self.train_samples

    def GetTestSamples(self):
        return self.test_samples

    def GetTestSamplesSampled(self, sampling_rate: float = 1.0, sampling_max_num: int = 1000000):
        """A sampled, formatted test batch for the periodic loss trace (lr2.py:216-230,428)."""
        if self.mode == "all":
            n = min(sampling_max_num, int(sampling_rate * len(self.test_samples)))
            idx = np.random.choice(len(self.test_samples), size=n, replace=False) if n else []
            return Sample.format_samples_sparse([self.test_samples[i] for i in idx])
        samples: List[Sample] = []
        for _ in range(int(sampling_max_num / self.batch_size)):
            try:
                samples.extend(self.NextBatchQueue(data_type="test"))
            except (DeadlineExceededError, QueueClosedError) as e:
                print("ignore exception: ", e)
                time.sleep(0.05)
        return Sample.format_samples_sparse(samples)

    def InitThreads(self, files, file_num, thread_num, sampling_rate=1.0, sess=None, is_train=True):
        q = self.train_data_queue if is_train else self.test_data_queue
        return [LoadDataThread(tid, files[tid:file_num:thread_num], sampling_rate, self.mode, q, is_train,
                               self.batch_size, self.coord,
                               None if self.seed is None else self.seed * 1000 + tid + (0 if is_train else 500))
                for tid in range(thread_num)]

    def LoadData(self) -> None:
        if self.train_threads is None:
            raise RuntimeError("DataProvider.init() must be called before LoadData()")
        if self.mode == "all":
            self.LoadDataAll(self.train_samples, self.train_threads)
            self.LoadDataAll(self.test_samples, self.test_threads)
        elif self.mode == "queue":
            log.info("start load train data queue ...")
            self.LoadDataQueue(self.train_threads)
            log.info("start load test data queue ...")
            self.LoadDataQueue(self.test_threads)

    def LoadDataAll(self, samples, threads) -> None:
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        for t in threads:
            samples.extend(t.samples)
        log.info("samples: %d" % len(samples))

    def LoadDataQueue(self, threads) -> None:
        for t in threads:
            t.start()

    def Shuffle(self) -> None:
        np.random.shuffle(self.train_samples)

    def NextBatch(self, data_type: str = "train"):
        if self.mode == "all":
            samples = self.train_samples if data_type == "train" else self.test_samples
            n = len(samples)
            for s in range(0, n, self.batch_size):
                yield samples[s:min(s + self.batch_size, n)]
        elif self.mode == "queue":
            yield self.NextBatchQueue(data_type)

    def NextBatchAll(self, data_type: str):
        samples = self.train_samples if data_type == "train" else self.test_samples
        n = len(samples)
        for s in range(0, n, self.batch_size):
            yield samples[s:min(s + self.batch_size, n)]

    def NextBatchQueue(self, data_type: str):
        b = self.train_data_batch if data_type == "train" else self.test_data_batch
        lines = b.next(timeout=self.dequeue_timeout)
        return [Sample().parse_line_libsvm(line) for line in lines]

    @staticmethod
    def GetFileList(input_files: str, num_workers: int, task_index: int) -> List[str]:
        file_list = [f for f in input_files.strip(",").split(",") if f.strip()]
        return [f.strip() for f in file_list[task_index:len(file_list):num_workers]]

    def close(self) -> None:
        """Queue teardown of lr2.py:471-475."""
        self.train_data_queue.close(cancel_pending_enqueues=True)
        self.test_data_queue.close(cancel_pending_enqueues=True)
        if self.coord is not None:
            self.coord.request_stop()
            try:
                self.coord.join(self.threads_dequeue, stop_grace_period_secs=5)
            except RuntimeError:
                pass
