"""ClusterSpec (R1): job name -> list of ``host:port`` tasks.

Same data as ``tf.train.ClusterSpec`` built from hard-coded lists (example.py:24-30) or from
``cluster_conf.json`` (cluster_conf.json:1-8, lr2.py:325-327).  On top of that it maps
tasks to ranks of the NVLink process group: worker k is rank k and also parameter-server
shard k; ``ps`` tasks own no tensors in GPU mode but still exist as roles.
"""
from __future__ import annotations

import json
from dataclasses import dataclass, field
from typing import Dict, List, Tuple


@dataclass(frozen=True)
class ClusterSpec:
    jobs: Dict[str, Tuple[str, ...]] = field(default_factory=dict)

    def __init__(self, spec):
        if isinstance(spec, ClusterSpec):
            jobs = dict(spec.jobs)
        else:
            jobs = {}
            for job, tasks in dict(spec).items():
                if isinstance(tasks, dict):                       # {index: address}
                    tasks = [tasks[i] for i in sorted(tasks)]
                tasks = tuple(str(t) for t in tasks)
                for t in tasks:
                    _split_hostport(t)
                jobs[str(job)] = tasks
        object.__setattr__(self, "jobs", jobs)

    # -- constructors ------------------------------------------------------------
    @classmethod
    def from_json(cls, path: str) -> "ClusterSpec":
        with open(path, "r") as f:
            return cls(json.load(f))

    @classmethod
    def local(cls, num_ps: int = 1, num_workers: int = 2, base_port: int = 2222) -> "ClusterSpec":
        """localhost cluster on consecutive ports, like example.py:24-26."""
        ps = [f"localhost:{base_port + i}" for i in range(num_ps)]
        wk = [f"localhost:{base_port + num_ps + i}" for i in range(num_workers)]
        return cls({"ps": ps, "worker": wk})

    # -- queries -----------------------------------------------------------------
    def as_dict(self) -> Dict[str, List[str]]:
        return {j: list(t) for j, t in self.jobs.items()}

    def job_tasks(self, job: str) -> List[str]:
        if job not in self.jobs:
            raise KeyError(f"no job {job!r} in cluster spec (jobs: {sorted(self.jobs)})")
        return list(self.jobs[job])

    def num_tasks(self, job: str) -> int:
        return len(self.jobs.get(job, ()))

    @property
    def num_workers(self) -> int:
        return self.num_tasks("worker")

    @property
    def num_ps(self) -> int:
        return self.num_tasks("ps")

    def task_address(self, job: str, task_index: int) -> str:
        tasks = self.job_tasks(job)
        if not 0 <= task_index < len(tasks):
            raise IndexError(f"task_index {task_index} out of range for job {job!r} ({len(tasks)} tasks)")
        return tasks[task_index]

    # -- process-group mapping ---------------------------------------------------
    def world_size(self) -> int:
        """Control-plane world: every ps task and every worker task is one process."""
        return self.num_ps + self.num_workers

    def global_rank(self, job: str, task_index: int) -> int:
        """Workers first (rank k = worker k = GPU k), then ps tasks."""
        self.task_address(job, task_index)
        return task_index if job == "worker" else self.num_workers + task_index

    def role_of_rank(self, rank: int) -> Tuple[str, int]:
        if rank < self.num_workers:
            return "worker", rank
        return "ps", rank - self.num_workers

    def rendezvous(self) -> Tuple[str, int]:
        """(host, port) of the control-plane store: the first ps task, else the first worker.
        ``localhost`` is pinned to 127.0.0.1 (container hostnames may not resolve)."""
        first = (self.jobs.get("ps") or self.jobs.get("worker"))[0]
        host, port = _split_hostport(first)
        return ("127.0.0.1" if host == "localhost" else host), port


def _split_hostport(addr: str) -> Tuple[str, int]:
    host, sep, port = addr.rpartition(":")
    if not sep or not host or not port.isdigit():
        raise ValueError(f"bad task address {addr!r} (want host:port)")
    return host, int(port)
