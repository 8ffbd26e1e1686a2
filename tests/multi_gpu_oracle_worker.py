"""Multi-GPU results checked against INDEPENDENT PyTorch oracles (float64, one process) — not against other b200ps
paths.  Run under torchrun, one rank per GPU (see test_gpu_multi.py).  Every rank regenerates every rank's batches
(the synthetic generators are pure functions of seed and offset), rank 0 evaluates the oracle on the host."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from b200ps import ops  # noqa: E402
from b200ps.parallel.context import init_from_env  # noqa: E402
from b200ps.ps.store import ParameterStore  # noqa: E402
from b200ps.train import MLPTrainer, SparseLRTrainer  # noqa: E402


def mlp_oracle(params, batches, lr, steps, counts=None):
    """`steps` synchronous steps in float64: every step applies the mean over the counted replicas of their
    per-replica mean gradients.  batches[r] = (x[T,B,D], y[T,B]); counts(t) -> replica ids aggregated at step t."""
    W1, B1, W2, B2 = (p.double().clone() for p in params)
    world = len(batches)
    losses = []
    for t in range(steps):
        who = list(range(world)) if counts is None else counts(t)
        acc = [torch.zeros_like(p) for p in (W1, B1, W2, B2)]
        for r in who:
            x, y = batches[r][0][t].double(), batches[r][1][t].long()
            ps = [p.clone().requires_grad_(True) for p in (W1, B1, W2, B2)]
            loss = torch.nn.functional.cross_entropy(torch.sigmoid(x @ ps[0] + ps[1]) @ ps[2] + ps[3], y)
            loss.backward()
            if r == 0:
                losses.append(loss.item())
            for a, p in zip(acc, ps):
                a += p.grad
        W1, B1, W2, B2 = (p - lr * a / len(who) for p, a in zip((W1, B1, W2, B2), acc))
    return (W1, B1, W2, B2), losses


def main():
    out_dir = sys.argv[1]
    ctx = init_from_env()
    dev, rank, world = ctx.device, ctx.rank, ctx.world
    res = {"world": world}
    names = ("weights/Variable", "biases/Variable", "weights/Variable_1", "biases/Variable_1")
    T, B = 12, 100

    def all_batches():
        out = []
        for r in range(world):
            x = torch.empty(T * B, 784, device=dev); y = torch.empty(T * B, device=dev, dtype=torch.int32)
            ops.synth_mnist(x, y, seed=7, offset=r * T * B)
            out.append((x.view(T, B, 784), y.view(T, B)))
        return out

    batches = all_batches()
    # ---- 1. synchronous MLP, all replicas aggregated (in-kernel push) == float64 big-batch SGD
    lr = 0.05
    tr = MLPTrainer(ctx, batch_size=B, lr=lr, mode="sync", use_graph=True)
    init = [tr.store.state_dict()[n].clone() for n in names]
    for t in range(T):
        tr.step(batches[rank][0][t], batches[rank][1][t])
    torch.cuda.synchronize(); ctx.control_barrier()
    got = tr.store.state_dict()
    if rank == 0:
        want, _ = mlp_oracle(init, [(x.cpu(), y.cpu()) for x, y in batches], lr, T)
        res["sync_mlp_max_abs_err"] = max(float((got[n].double() - w).abs().max()) for n, w in zip(names, want))
        res["sync_mlp_global_step"] = tr.global_step
    ctx.control_barrier()

    # ---- 2. replicas_to_aggregate = world - 1: the GPU path aggregates the rotating set documented in DESIGN.md §4
    #         (replica r counts at push e = 1, 2, ... iff (r - e mod world) mod world < R); oracle applies exactly that
    R = world - 1
    tr_r = MLPTrainer(ctx, batch_size=B, lr=lr, mode="sync", use_graph=False, replicas_to_aggregate=R)
    init_r = [tr_r.store.state_dict()[n].clone() for n in names]
    for t in range(T):
        tr_r.step(batches[rank][0][t], batches[rank][1][t])
    torch.cuda.synchronize(); ctx.control_barrier()
    got_r = tr_r.store.state_dict()
    if rank == 0:
        counts = lambda t: [r for r in range(world) if (r - (t + 1) % world + world) % world < R]   # noqa: E731
        want, _ = mlp_oracle(init_r, [(x.cpu(), y.cpu()) for x, y in batches], lr, T, counts)
        res["sync_backup_max_abs_err"] = max(float((got_r[n].double() - w).abs().max()) for n, w in zip(names, want))
        res["sync_backup_global_step"] = tr_r.global_step
    ctx.control_barrier()

    # ---- 3. asynchronous push conserves every update: K pushes of the constant gradient (rank + 1) from every rank
    #         leave master = init - lr * K * sum(rank + 1) exactly (all values are small integers times lr = 2^-6)
    st = ParameterStore(ctx.pool)
    st.register("v", (3, 1000), init="zeros")
    st.build(); st.initialize()
    K = 25
    st.grad("v").fill_(float(rank + 1))
    torch.cuda.synchronize(); ctx.control_barrier()
    for _ in range(K):
        st.push_async(2.0 ** -6)
    torch.cuda.synchronize(); ctx.control_barrier()
    v = st.state_dict()["v"]
    res["async_conserved"] = bool((v == -(2.0 ** -6) * K * world * (world + 1) / 2).all())
    res["async_global_step"] = st.global_step
    ctx.control_barrier()

    # ---- 4. sparse LR, synchronous: weight VALUES against the float64 dense restatement — the direct kernel on
    #         uniform ids, then the unique-id step (sparse_dedup.cu) on ids that repeat inside and across the batches
    from b200ps.models.sparse_lr import sparse_lr_reference
    F, SB, nnz, spread, ST = 1 << 16, 200, 30, 6, 5
    for tag, dedupe in (("sparse_sync", False), ("sparse_unique_sync", True)):
        ts = SparseLRTrainer(ctx, features=F, lr=0.5, mode="sync", use_graph=False, dedupe=dedupe)
        W0 = ts.store.state_dict()
        w_name, b_name = ts.model.W, ts.model.Bn
        sb = []
        for r in range(world):
            rp = torch.empty(SB + 1, device=dev, dtype=torch.int32); fid = torch.empty(SB * (nnz + spread), device=dev, dtype=torch.int64)
            fv = torch.empty(SB * (nnz + spread), device=dev); yy = torch.empty(SB, device=dev)
            ops.synth_sparse(rp, fid, fv, yy, nnz_mean=nnz, spread=spread, features=F, seed=11, offset=r * SB)
            if dedupe:                           # F * u^8: a handful of ids carries most non-zeros, on every rank
                g = torch.Generator(device=dev).manual_seed(40 + r)
                u = torch.rand(fid.numel(), device=dev, generator=g, dtype=torch.float64)
                fid.copy_((u.pow(8) * F).long().clamp_(0, F - 1))
            sb.append((rp, fid, fv, yy))
        for _ in range(ST):
            ts.step(sb[rank])
        torch.cuda.synchronize(); ctx.control_barrier()
        got_s = ts.store.state_dict()
        if rank == 0:
            W = W0[w_name].double().reshape(-1).clone(); b = W0[b_name].double().reshape(()).clone()
            for _ in range(ST):
                dW = torch.zeros_like(W); db = 0.0
                for r in range(world):
                    rp, fid, fv, yy = (t.cpu() for t in sb[r])
                    _, _, (ids, vals), g_b = sparse_lr_reference(W, b, rp, fid, fv, yy)
                    dW.index_add_(0, ids, vals); db = db + g_b
                W -= 0.5 / world * dW; b = b - 0.5 / world * db
            res[tag + "_max_abs_err"] = float((got_s[w_name].double().reshape(-1) - W).abs().max())
            res[tag + "_bias_err"] = float((got_s[b_name].double().reshape(()) - b).abs())
            res[tag + "_global_step"] = ts.global_step
            if dedupe:
                d, nz = ts.model.last_unique_counts(sb[rank][1].numel(), 0)
                res["sparse_unique_distinct_ratio"] = d / max(nz, 1)
        ctx.control_barrier()
    # ---- 5. large dense groups: the store's NVLS path (multimem.ld_reduce / multimem.st through the switch) against
    #         a torch restatement of the synchronous update, fp32 and bf16 (+ fp32 master), then the collective pull
    import torch.distributed as dist
    for tag, dt, n_el in (("f32", "fp32", 3 << 19), ("bf16", "bf16:0", 1 << 21)):
        sn = ParameterStore(ctx.pool)
        sn.register("big", (n_el // 1024, 1024), group=dt, init="normal", stddev=0.5)
        sn.build(); sn.initialize()
        g = sn.groups[dt]
        res[f"nvls_{tag}_used"] = bool(sn._nvls(g))
        w0 = sn.state_dict()["big"].reshape(-1).clone()
        ref = w0.double()
        errs = []
        for t in range(3):
            gen = torch.Generator(device=dev).manual_seed(1000 * t + rank)
            mine = torch.randn(n_el, device=dev, generator=gen) * 0.1
            sn.grad("big").reshape(-1).copy_(mine)
            allg = []
            for r in range(world):          # every rank regenerates every rank's gradient (as stored: bf16-rounded for bf16)
                gr = torch.Generator(device=dev).manual_seed(1000 * t + r)
                v = torch.randn(n_el, device=dev, generator=gr) * 0.1
                allg.append(v.to(sn.grad("big").dtype).double().cpu())
            torch.cuda.synchronize(); ctx.control_barrier()
            sn.push_sync(0.25)
            torch.cuda.synchronize(); ctx.control_barrier()
            ref = ref - 0.25 / world * torch.stack(allg).sum(0)
            errs.append(float((sn.state_dict()["big"].reshape(-1).double() - ref).abs().max()))
        work = sn.var("big").reshape(-1).double().cpu()
        res[f"nvls_{tag}_master_err"] = max(errs)
        res[f"nvls_{tag}_work_err"] = float((work - ref).abs().max())        # the broadcast reached this replica
        res[f"nvls_{tag}_global_step"] = sn.global_step
        lo_own, hi_own = g.layout.bounds[rank], g.layout.bounds[rank + 1]
        flat = sn.flat_work(dt)
        if g.master is None:                # fp32 group: the owned slice of the working copy IS the master — keep it
            flat[:lo_own].zero_(); flat[hi_own:].zero_()
        else:
            flat.zero_()                    # collective pull restores everything from the owners' fp32 masters
        torch.cuda.synchronize(); ctx.control_barrier()
        sn.pull(collective=True)
        torch.cuda.synchronize(); ctx.control_barrier()
        res[f"nvls_{tag}_pull_err"] = float((sn.var("big").reshape(-1).double().cpu() - ref).abs().max())
        allw = [None] * world
        dist.all_gather_object(allw, float(sn.var("big").double().sum().item()), group=ctx.group)
        res[f"nvls_{tag}_replicas_identical"] = len(set(allw)) == 1
        del sn
        ctx.control_barrier()
    # ---- 6. sparse wide & deep, synchronous, owner-side rows (request / serve / loss / fold rounds): table VALUES
    #         after 3 steps against the float64 oracle (mean of the replicas' gradients per step)
    from b200ps.models.wide_deep import wide_deep_reference
    from b200ps.train import SparseWideDeepTrainer
    FW, BW = 1 << 13, 160
    tw = SparseWideDeepTrainer(ctx, features=FW, dim=128, lr=0.5, mode="sync", use_graph=False)
    res["wd_owner_step"] = bool(tw.owner_step)
    mw = tw.model
    w_init = {k: v.clone() for k, v in tw.store.state_dict().items()}
    wb = []
    for r in range(world):
        rp = torch.empty(BW + 1, device=dev, dtype=torch.int32); fid = torch.empty(BW * 40, device=dev, dtype=torch.int64)
        fv = torch.empty(BW * 40, device=dev); yy = torch.empty(BW, device=dev)
        ops.synth_sparse(rp, fid, fv, yy, nnz_mean=32, spread=8, features=FW, seed=17, offset=r * BW)
        wb.append((rp, fid, fv, yy))
    for _ in range(3):
        tw.step(wb[rank])
    torch.cuda.synchronize(); ctx.control_barrier()
    got_w = tw.store.state_dict()
    if rank == 0:
        W = w_init[mw.W].double().reshape(-1); E = w_init[mw.E].double(); w2v = w_init[mw.W2].double().reshape(-1)
        bb = w_init[mw.Bn].double().reshape(())
        cpu = [tuple(t.cpu() for t in b) for b in wb]
        for _ in range(3):
            acc = None
            for r in range(world):
                _, _, gr = wide_deep_reference(W, E, w2v, bb, *cpu[r])
                acc = gr if acc is None else {k: acc[k] + gr[k] for k in gr}
            W = W - 0.5 / world * acc["W"]; E = E - 0.5 / world * acc["E"]
            w2v = w2v - 0.5 / world * acc["w2"]; bb = bb - 0.5 / world * acc["b"]
        res["wd_sync_E_err"] = float((got_w[mw.E].double() - E).abs().max())
        res["wd_sync_W_err"] = float((got_w[mw.W].double().reshape(-1) - W).abs().max())
        res["wd_sync_w2_err"] = float((got_w[mw.W2].double().reshape(-1) - w2v).abs().max())
        res["wd_sync_b_err"] = float((got_w[mw.Bn].double().reshape(()) - bb).abs())
        res["wd_sync_global_step"] = tw.global_step
    ctx.control_barrier()
    del tw
    if rank == 0:
        json.dump(res, open(os.path.join(out_dir, "oracle_result.json"), "w"))
    print("rank", rank, json.dumps(res), flush=True)
    del tr, tr_r, st, ts
    ctx.shutdown()


if __name__ == "__main__":
    main()
