"""Control-plane liveness check on the CPU (gloo): rank 1 dies, rank 0 raises PeerLost and abandons the job."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ps.parallel.context import PeerLost, init_from_env  # noqa: E402


def main():
    out = sys.argv[1]
    ctx = init_from_env(device="cpu")
    res = {"checks_ok": 0}
    for step in range(20):
        if ctx.rank == ctx.world - 1 and step == 3:
            os._exit(0)
        t0 = time.perf_counter()
        try:
            ctx.check_peers(timeout_s=3.0)
            res["checks_ok"] += 1
        except PeerLost as e:
            res.update(lost=True, seconds=time.perf_counter() - t0, message=str(e))
            break
    if ctx.rank == 0:
        json.dump(res, open(out, "w"))
    ctx.abandon()


if __name__ == "__main__":
    main()
