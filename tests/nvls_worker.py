"""torchrun worker: K1-NVLS (multimem push) against a plain PyTorch restatement, plus timing vs NCCL."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    out_dir = sys.argv[1]
    dist.init_process_group("nccl")
    rank, world = dist.get_rank(), dist.get_world_size()
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dev = torch.device("cuda", torch.cuda.current_device())
    res = {"world": world}
    try:
        from b200ps.parallel.nvls import NvlsGroup
        n, lr = 1 << 20, 0.5
        g0 = torch.Generator(device=dev).manual_seed(7)
        init = torch.randn(n, device=dev, generator=g0)                       # same on every rank
        grp = NvlsGroup(n, lr, init=init)
        ref = init.clone()
        ok = True
        for step in range(3):
            gr = torch.Generator(device=dev).manual_seed(100 * step + rank)
            mine = torch.randn(n, device=dev, generator=gr)
            grp.grads[:n].copy_(mine)
            allg = [torch.empty_like(mine) for _ in range(world)]
            dist.all_gather(allg, mine)
            ref -= lr / world * torch.stack(allg).sum(0)
            grp.push_sync()
            torch.cuda.synchronize()
            err = (grp.work[:n] - ref).abs().max().item()
            ok = ok and err < 1e-5
            res["err_step%d" % step] = err
        res["ok"] = bool(ok)
        res["global_step"] = grp.global_step
        digests = [None] * world
        dist.all_gather_object(digests, float(grp.work[:n].double().sum().item()))
        res["replicas_identical"] = len(set(digests)) == 1
        # timing at 64 MB: multimem push (incl. its two barriers) vs NCCL reduce_scatter + add + all_gather
        n2 = 1 << 24
        big = NvlsGroup(n2, lr)
        big.grads.normal_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(3):
            big.push_sync()
        torch.cuda.synchronize(); dist.barrier(); e0.record()
        for _ in range(10):
            big.push_sync()
        e1.record(); torch.cuda.synchronize()
        res["nvls_push_us_64MB"] = e0.elapsed_time(e1) * 100.0
        gbuf = torch.randn(n2, device=dev); wbuf = torch.zeros(n2, device=dev)
        shard = torch.empty(n2 // world, device=dev)

        def nccl_push():
            dist.reduce_scatter_tensor(shard, gbuf)
            wbuf[rank * (n2 // world):(rank + 1) * (n2 // world)].add_(shard, alpha=-lr / world)
            dist.all_gather_into_tensor(wbuf, wbuf[rank * (n2 // world):(rank + 1) * (n2 // world)].clone())
        for _ in range(3):
            nccl_push()
        torch.cuda.synchronize(); dist.barrier(); e0.record()
        for _ in range(10):
            nccl_push()
        e1.record(); torch.cuda.synchronize()
        res["nccl_push_us_64MB"] = e0.elapsed_time(e1) * 100.0
    except Exception as e:  # noqa: BLE001
        import traceback
        res["error"] = f"{type(e).__name__}: {e}"[:500]
        res["trace"] = traceback.format_exc()[-800:]
    if rank == 0:
        json.dump(res, open(os.path.join(out_dir, "nvls_result.json"), "w"))
        print("NVLS_RESULT", res, flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
