"""Race-check substitute for the NVLink backend: thousands of synchronous in-kernel-push steps with a
pseudo-random delay injected into every CTA of every rank (b200ps_debug_set_skew_ns), on the CHECKED library
(device-side assertions: packet slots written once per epoch, epochs only move forward, owners in range).  The
skewed run must reproduce the unskewed run bit for bit and keep the replicas identical."""
import hashlib
import json
import os
import sys

os.environ["B200PS_CHECKED"] = "1"          # before b200ps.ops.native is imported

import torch  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from b200ps import ops  # noqa: E402
from b200ps.ops import native  # noqa: E402
from b200ps.parallel.context import init_from_env  # noqa: E402
from b200ps.train import MLPTrainer  # noqa: E402


def digest(t):
    return hashlib.sha1(t.detach().contiguous().view(torch.uint8).cpu().numpy().tobytes()).hexdigest()


def main():
    out_dir, steps, skew_ns = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    ctx = init_from_env()
    dev, rank = ctx.device, ctx.rank
    lib = native.load()
    assert native.lib_path().endswith("libb200ps_checked.so")
    data = torch.empty(4000, 784, device=dev); labels = torch.empty(4000, device=dev, dtype=torch.int32)
    ops.synth_mnist(data, labels, seed=3, offset=rank * 4000)
    out = {}
    for name, skew in (("plain", 0), ("skewed", skew_ns)):
        tr = MLPTrainer(ctx, batch_size=100, lr=0.002, mode="sync", use_graph=True)
        assert tr.fused_push
        lib.b200ps_debug_set_skew_ns(skew)
        tr.step(data[:100], labels[:100]); tr.step(data[:100], labels[:100]); tr.step(data[:100], labels[:100])
        tr.attach_dataset(data, labels, steps_per_graph=8)
        tr.run_resident(steps)
        torch.cuda.synchronize()
        lib.b200ps_debug_set_skew_ns(0)
        ctx.control_barrier()
        import torch.distributed as dist
        ds = [None] * ctx.world
        dist.all_gather_object(ds, digest(tr.store.flat_work("fp32")), group=ctx.group)
        out[name] = {"digest": ds[0], "replicas_identical": len(set(ds)) == 1, "global_step": tr.global_step,
                     "loss": tr.read_loss()}
        del tr
        ctx.control_barrier()
    out["bitwise_equal"] = out["plain"]["digest"] == out["skewed"]["digest"]
    if rank == 0:
        json.dump(out, open(os.path.join(out_dir, "stress_result.json"), "w"))
    print("rank", rank, json.dumps(out), flush=True)
    ctx.shutdown()


if __name__ == "__main__":
    main()
