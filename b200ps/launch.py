"""Local multi-process launcher (P30): the reference's "start one process per role by hand on
localhost" recipe (README.md:11-16, example.py:8-11), automated.

    python -m b200ps.launch --num_ps 1 --num_workers 2 example -- --sync --training_epochs 1
    python -m b200ps.launch --num_ps 0 --num_workers 8 lr2 -- --backend nvlink --train synthetic ...

It writes a ``cluster_conf.json`` (the file lr2 reads, cluster_conf.json:1-8) on free
localhost ports, spawns one ``python -m b200ps.apps.<app>`` per task with ``--job_name`` /
``--task_index``, streams their output with a ``[job:task]`` prefix and returns the first
non-zero exit code.
"""
from __future__ import annotations

import argparse
import json
import os
import socket
import subprocess
import sys
import threading


def free_ports(n: int):
    socks, ports = [], []
    for _ in range(n):
        s = socket.socket()
        s.bind(("127.0.0.1", 0))
        socks.append(s)
        ports.append(s.getsockname()[1])
    for s in socks:
        s.close()
    return ports


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--num_ps", type=int, default=1)
    ap.add_argument("--num_workers", type=int, default=2)
    ap.add_argument("--workdir", default=".")
    ap.add_argument("--keep_going", action="store_true", help="let surviving workers finish after a worker failure")
    ap.add_argument("--timeout", type=float, default=0, help="kill everything after this many seconds (0 = never)")
    ap.add_argument("app", choices=["example", "lr2"])
    ap.add_argument("app_args", nargs=argparse.REMAINDER)
    a = ap.parse_args(argv)
    extra = [x for x in a.app_args if x != "--"]
    ports = free_ports(a.num_ps + a.num_workers)
    spec = {"worker": [f"localhost:{p}" for p in ports[a.num_ps:]]}
    if a.num_ps:
        spec["ps"] = [f"localhost:{p}" for p in ports[:a.num_ps]]
    os.makedirs(a.workdir, exist_ok=True)
    conf = os.path.join(os.path.abspath(a.workdir), "cluster_conf.json")
    with open(conf, "w") as f:
        json.dump(spec, f, indent=4)
    env = dict(os.environ)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env["PYTHONPATH"] = root + os.pathsep + env.get("PYTHONPATH", "")
    # Every task is a torch process whose CPU math is tiny (the GPU backend does none at all): left alone, each
    # spins one intra-op thread per core and the tasks fight for them (measured with the gloo backend, 1 ps + 2
    # workers on 8 cores: 54 ms/step, 9 ms/step with one thread per task).  Same default as torchrun.
    if "OMP_NUM_THREADS" not in env:
        env["OMP_NUM_THREADS"] = str(max(1, (os.cpu_count() or 1) // max(1, 2 * (a.num_ps + a.num_workers))))
    procs = []
    for job, n in (("ps", a.num_ps), ("worker", a.num_workers)):
        for i in range(n):
            cmd = [sys.executable, "-m", f"b200ps.apps.{a.app}", f"--job_name={job}", f"--task_index={i}"]
            if a.app == "example":
                cmd += ["--ps_hosts", ",".join(spec.get("ps", [])), "--worker_hosts", ",".join(spec["worker"])]
            else:
                cmd += ["--cluster_conf", conf]
            p = subprocess.Popen(cmd + extra, cwd=a.workdir, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                                 text=True)
            procs.append((f"{job}:{i}", p))

    def pump(tag, p):
        for line in p.stdout:
            sys.stdout.write(f"[{tag}] {line}")
            sys.stdout.flush()

    threads = [threading.Thread(target=pump, args=tp, daemon=True) for tp in procs]
    for t in threads:
        t.start()
    rc = 0
    workers = [(t, p) for t, p in procs if t.startswith("worker")]
    servers = [(t, p) for t, p in procs if t.startswith("ps")]
    try:
        for tag, p in workers:
            try:
                code = p.wait(timeout=a.timeout or None)
            except subprocess.TimeoutExpired:
                code = -9
            if code != 0 and rc == 0:
                rc = code
                if not a.keep_going:        # one worker failed: do not leave the others hanging
                    for _, q in procs:
                        if q.poll() is None:
                            q.terminate()
        for tag, p in servers:              # ps tasks leave join() once every done-token is in;
            try:                            # after a worker crash they never will: stop them
                code = p.wait(timeout=10 if rc == 0 else 1)
            except subprocess.TimeoutExpired:
                p.terminate()
                code = 0 if rc != 0 else -9
            if code != 0 and rc == 0:
                rc = code
    finally:
        for _, p in procs:
            if p.poll() is None:
                p.kill()
    for t in threads:
        t.join(timeout=2)
    return rc


if __name__ == "__main__":
    sys.exit(main())
