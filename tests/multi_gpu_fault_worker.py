"""Fault injection on the NVLink backend: rank 1 dies between two synchronous steps; rank 0 finds out
through the control-plane liveness check BEFORE it launches a step that would wait on the dead rank's packets,
raises PeerLost within the timeout, abandons the job without barriers and exits normally."""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from b200ps import ops  # noqa: E402
from b200ps.parallel.context import PeerLost, init_from_env  # noqa: E402
from b200ps.train import MLPTrainer  # noqa: E402


def main():
    out_dir, die_at = sys.argv[1], int(sys.argv[2])
    ctx = init_from_env()
    dev, rank = ctx.device, ctx.rank
    tr = MLPTrainer(ctx, batch_size=100, lr=0.01, mode="sync", use_graph=True)
    x = torch.empty(100, 784, device=dev); y = torch.empty(100, device=dev, dtype=torch.int32)
    ops.synth_mnist(x, y, seed=1, offset=rank * 100)
    res = {"steps_done": 0}
    t_lost = None
    for step in range(50):
        if rank == ctx.world - 1 and step == die_at:
            torch.cuda.synchronize()
            os._exit(0)              # no teardown, no goodbye: the shard is gone (exit code 0 keeps torchrun quiet)
        t0 = time.perf_counter()
        try:
            ctx.check_peers(timeout_s=4.0)
        except PeerLost as e:
            t_lost = time.perf_counter() - t0
            res.update(lost=True, message=str(e), detect_seconds=t_lost, global_step=tr.global_step)
            break
        tr.step(x, y)
        torch.cuda.synchronize()
        res["steps_done"] = step + 1
    if rank == 0:
        json.dump(res, open(os.path.join(out_dir, "fault_result.json"), "w"))
    print("rank", rank, json.dumps(res), flush=True)
    del tr
    ctx.abandon()


if __name__ == "__main__":
    main()
