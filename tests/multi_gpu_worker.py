"""Body of the multi-GPU tests: run under torchrun, one rank per GPU (see test_gpu_multi.py)."""
import hashlib
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from b200ps import ops  # noqa: E402
from b200ps.io.saver import Saver, store_variables  # noqa: E402
from b200ps.parallel.context import init_from_env  # noqa: E402
from b200ps.ps.store import ParameterStore  # noqa: E402
from b200ps.train import MLPTrainer, SparseLRTrainer, WideMLPTrainer  # noqa: E402


def digest(t: torch.Tensor) -> str:
    return hashlib.sha1(t.detach().contiguous().view(torch.uint8).cpu().numpy().tobytes()).hexdigest()


def gather(ctx, obj):
    import torch.distributed as dist
    out = [None] * ctx.world
    dist.all_gather_object(out, obj, group=ctx.group)
    return out


def main():
    out_dir = sys.argv[1]
    ctx = init_from_env()
    dev, rank, world = ctx.device, ctx.rank, ctx.world
    res = {"world": world}

    # ---- sync MLP: replicas stay bit-identical, one global step per aggregate, loss falls
    tr = MLPTrainer(ctx, batch_size=100, lr=0.01, mode="sync", use_graph=True)
    x = torch.empty(20, 100, 784, device=dev); y = torch.empty(20, 100, device=dev, dtype=torch.int32)
    ops.synth_mnist(x.view(-1, 784), y.view(-1), seed=1, offset=rank * 2000)
    losses = []
    for i in range(60):
        tr.step(x[i % 20], y[i % 20])
        if i in (0, 59):
            losses.append(tr.read_loss())
    torch.cuda.synchronize()
    d = digest(tr.store.flat_work("fp32"))
    res["sync_digests_equal"] = len(set(gather(ctx, d))) == 1
    res["sync_global_step"] = tr.global_step
    res["sync_loss"] = losses
    if rank == 0:
        Saver().save(store_variables(tr.store), os.path.join(out_dir, "model.ckpt"), tr.global_step)
    acc, _ = tr.evaluate(x.view(-1, 784), y.view(-1))
    res["sync_acc"] = acc
    ctx.control_barrier()
    # the same 60 steps with the push as a separate kernel (K13 gradients + push_sync_ll) and with the
    # four-kernel tcgen05 chain: the in-kernel push must reproduce the former bit for bit
    res["sync_fused_push_used"] = bool(tr.fused_push)
    os.environ["B200PS_FUSED_PUSH"] = "0"
    tr_sep = MLPTrainer(ctx, batch_size=100, lr=0.01, mode="sync", use_graph=True)
    os.environ.pop("B200PS_FUSED_PUSH")
    tr_chain = MLPTrainer(ctx, batch_size=100, lr=0.01, mode="sync", use_graph=True, fused_step=False)
    for i in range(60):
        tr_sep.step(x[i % 20], y[i % 20])
        tr_chain.step(x[i % 20], y[i % 20])
    torch.cuda.synchronize()
    res["sync_fused_push_bitwise"] = (not tr_sep.fused_push) and digest(tr_sep.store.flat_work("fp32")) == d
    res["sync_chain_close"] = float((tr_chain.store.flat_work("fp32") - tr.store.flat_work("fp32")).abs().max())
    res["sync_steps_all"] = [tr_sep.global_step, tr_chain.global_step]
    ctx.control_barrier()

    # ---- async MLP (round-robin placement): the chief's counter sees every worker's update
    tr2 = MLPTrainer(ctx, batch_size=100, lr=0.01, mode="async", use_graph=True, placement="round_robin_variable")
    res["async_in_kernel"] = bool(tr2.fused_async)
    l2 = []
    for i in range(40):
        tr2.step(x[i % 20], y[i % 20])
        if i in (0, 39):
            l2.append(tr2.read_loss())
    torch.cuda.synchronize()
    ctx.control_barrier()
    res["async_global_step"] = tr2.global_step
    res["async_loss"] = l2

    # ---- wide MLP (bf16 group + fp32 group) at a small size
    tw = WideMLPTrainer(ctx, batch_size=256, layers=3, hidden=512, lr=0.5, use_graph=True, fused_pull=True)
    xb = torch.empty(256, 512, device=dev, dtype=torch.bfloat16); yb = torch.empty(256, device=dev, dtype=torch.int32)
    ops.synth_dense_bf16(xb, yb, classes=512, seed=3, offset=rank * 256)
    lw = []
    for i in range(12):
        tw.step(xb, yb)
        if i in (0, 11):
            lw.append(tw.read_loss())
    torch.cuda.synchronize()
    tw.store.pull()                                   # refresh every replica from the owners' masters
    torch.cuda.synchronize()
    ctx.control_barrier()
    res["wide_fused_pull"] = bool(tw.fused_pull)
    res["wide_digests_equal"] = len(set(gather(ctx, digest(tw.store.flat_work("bf16:1"))))) == 1
    res["wide_loss"] = lw
    ctx.control_barrier()

    # ---- sparse LR, async then sync
    F, B, nnz, spread = 1 << 20, 500, 50, 10
    for mode in ("async", "sync"):
        ts = SparseLRTrainer(ctx, features=F, lr=0.5, mode=mode, use_graph=False)
        rp = torch.empty(B + 1, device=dev, dtype=torch.int32); fid = torch.empty(B * (nnz + spread), device=dev, dtype=torch.int64)
        fv = torch.empty(B * (nnz + spread), device=dev); yy = torch.empty(B, device=dev)
        ops.synth_sparse(rp, fid, fv, yy, nnz_mean=nnz, spread=spread, features=F, seed=5, offset=rank * B)
        for i in range(10):
            ts.step((rp, fid, fv, yy))
        torch.cuda.synchronize()
        ctx.control_barrier()
        res[f"sparse_{mode}_global_step"] = ts.global_step
        res[f"sparse_{mode}_loss"] = ts.read_loss()
        ctx.control_barrier()

    # ---- stand-alone barrier kernel
    st = ParameterStore(ctx.pool).build()
    for _ in range(5):
        st.barrier()
    torch.cuda.synchronize()
    res["barrier_epoch"] = int(st.pads.view(rank, torch.int32)[97].item())
    ctx.control_barrier()
    if rank == 0:
        json.dump(res, open(os.path.join(out_dir, "result.json"), "w"))
    print("rank", rank, json.dumps(res), flush=True)
    ctx.control_barrier()
    del tr, tr_sep, tr_chain, tr2, tw, ts, st
    ctx.shutdown()


if __name__ == "__main__":
    main()
