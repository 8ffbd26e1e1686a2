"""BASELINE.json config 5: parameter-server push/pull sweep, 1 KB - 1 GB, b200ps kernels vs NCCL.

Run under torchrun (one rank per GPU).  For every size S (bytes of fp32 parameters) it times

    push_sync : the ParameterStore's push (reduce-scatter + mean + SGD + broadcast, flags included): LL packets for
                small groups, K1 over peer loads/stores, and from 1 MB K1-NVLS (multimem.ld_reduce / multimem.st) —
                both timed where both apply — vs NCCL reduce_scatter(AVG) + add_ + all_gather   (same result)
    pull      : reader-driven pull kernel (all-gather of owner slices) and the collective owner-side multimem.st
                broadcast vs NCCL all_gather_into_tensor
    push_async: b200ps K3 (peer red.add of -lr*g into the owners, no barrier)
                vs NCCL reduce_scatter(SUM) + add_                        (closest library form)

and reports microseconds (CUDA events, max over ranks) plus the per-GPU NVLink payload rate:
push_sync moves 2*(N-1)/N*S bytes per direction per GPU, pull and push_async (N-1)/N*S.
Roofline denominator: 770 GB/s measured peer copy per direction (B200_PROFILING.md).
"""
import datetime
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ps.parallel.context import init_from_env  # noqa: E402
from b200ps.ps.store import ParameterStore  # noqa: E402

LINK_GBS = 770.0


def timeit(fn, iters, ctx):
    for _ in range(3):
        fn()
    ctx.control_barrier()
    torch.cuda.synchronize()
    st, en = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    st.record()
    for _ in range(iters):
        fn()
    en.record()
    torch.cuda.synchronize()
    return ctx.all_reduce_max(st.elapsed_time(en) / iters * 1e3)


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", timeout=datetime.timedelta(seconds=300), device_id=torch.device(f"cuda:{local}"))
    ctx = init_from_env()                       # adds the gloo control group next to NCCL
    N, dev = ctx.world, ctx.device
    sizes = [1 << k for k in range(10, 31, 2)]  # 1 KB .. 1 GB
    max_bytes = int(os.environ.get("SWEEP_MAX_BYTES", str(1 << 30)))
    rows = []
    for S in sizes:
        if S > max_bytes:
            break
        n = S // 4
        store = ParameterStore(ctx.pool)
        store.register("p", (n,), init="zeros")
        store.build()
        store.initialize()
        store.flat_grads("fp32").normal_()
        iters = 200 if S <= (1 << 20) else (30 if S <= (1 << 26) else 8)
        g0 = store.groups["fp32"]
        has_mc = store._nvls(g0, True)             # >= 1 MB groups live in multicast memory (when the box has NVLS)
        has_nvls = store._nvls(g0)                 # ... and push through the switch by default up to 512 MB
        t_push = timeit(lambda: store.push_sync(1e-3), iters, ctx)          # the store's default path for this size
        t_push_p2p = t_push_nvls = t_pull_nvls = None
        if has_mc:
            t_push_nvls = timeit(lambda: store.push_sync(1e-3, nvls=True), iters, ctx)
            t_push_p2p = timeit(lambda: store.push_sync(1e-3, nvls=False), iters, ctx)   # K1 over peer loads / stores
            t_pull_nvls = timeit(lambda: store.pull(collective=True), iters, ctx)   # owner-side multimem.st broadcast
        t_pull = timeit(lambda: store.pull(include_self=False), iters, ctx)
        t_async = timeit(lambda: store.push_async(1e-3, count_step=False), iters, ctx)
        # NCCL twin on plain tensors of the same padded size
        total = store.groups["fp32"].layout.total
        shard = total // N
        w = torch.zeros(total, device=dev); g = torch.randn(total, device=dev); sl = torch.zeros(shard, device=dev)
        master = torch.zeros(shard, device=dev)

        def nccl_push():
            dist.reduce_scatter_tensor(sl, g, op=dist.ReduceOp.AVG)
            master.add_(sl, alpha=-1e-3)
            dist.all_gather_into_tensor(w, master)

        def nccl_async():
            dist.reduce_scatter_tensor(sl, g, op=dist.ReduceOp.SUM)
            master.add_(sl, alpha=-1e-3)

        n_push = timeit(nccl_push, iters, ctx)
        n_pull = timeit(lambda: dist.all_gather_into_tensor(w, master), iters, ctx)
        n_async = timeit(nccl_async, iters, ctx)
        f = (N - 1) / N
        rec = {"bytes": S, "n_gpus": N,
               "push_sync_us": round(t_push, 2), "nccl_push_us": round(n_push, 2),
               "push_sync_path": "nvls" if has_nvls else ("ll" if g0.t_ll_grads is not None and store.use_ll else "p2p"),
               "push_sync_p2p_us": None if t_push_p2p is None else round(t_push_p2p, 2),
               "push_sync_nvls_us": None if t_push_nvls is None else round(t_push_nvls, 2),
               "pull_nvls_us": None if t_pull_nvls is None else round(t_pull_nvls, 2),
               "push_vs_nccl": round(n_push / t_push, 3),
               "pull_best_vs_nccl": round(n_pull / min(t_pull, t_pull_nvls or t_pull), 3),
               "pull_us": round(t_pull, 2), "nccl_allgather_us": round(n_pull, 2),
               "push_async_us": round(t_async, 2), "nccl_rs_apply_us": round(n_async, 2),
               "push_sync_GBs_per_dir": round(2 * f * S / t_push / 1e3, 1),
               "pull_GBs_per_dir": round(f * S / t_pull / 1e3, 1),
               "push_async_GBs_per_dir": round(f * S / t_async / 1e3, 1),
               "push_sync_frac_of_link": round(2 * f * S / t_push / 1e3 / LINK_GBS, 3),
               "pull_frac_of_link": round(f * S / t_pull / 1e3 / LINK_GBS, 3)}
        rows.append(rec)
        if ctx.rank == 0:
            print(json.dumps(rec), flush=True)
        del store, w, g, sl, master
        torch.cuda.empty_cache()
    if ctx.rank == 0:
        os.makedirs("gpurun_out", exist_ok=True)
        json.dump(rows, open(f"gpurun_out/pushpull_sweep_n{N}.json", "w"), indent=1)
    ctx.shutdown()


if __name__ == "__main__":
    main()
