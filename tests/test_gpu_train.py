"""End-to-end GPU training paths: trainers, ParameterStore, virtual-rank sync/async."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _mnist_batches(n, B=100, seed=1, offset=0):
    from b200ps import ops
    x = torch.empty(n * B, 784, device=DEV); y = torch.empty(n * B, device=DEV, dtype=torch.int32)
    ops.synth_mnist(x, y, seed=seed, offset=offset)
    return x.view(n, B, 784), y.view(n, B)


@pytest.mark.parametrize("mode,fused", [("sync", True), ("sync", False), ("async", True), ("async", False)])
def test_mlp_step_matches_reference_math(native_lib, mode, fused):
    """One step of every single-replica path — one-launch K13 (in-place SGD / in-kernel Hogwild) and the
    tcgen05 chain with a separate push — against the fp32 restatement of example.py:95-118."""
    from b200ps.models.mlp import mlp_reference_step
    from b200ps.parallel.context import init_from_env
    from b200ps.train import MLPTrainer
    ctx = init_from_env(virtual_world=1)
    tr = MLPTrainer(ctx, batch_size=100, lr=0.0005, mode=mode, use_graph=False, fused_step=fused)
    assert tr.fused == fused and tr.fused_apply == (fused and mode == "sync") and tr.fused_async == (fused and mode == "async")
    xs, ys = _mnist_batches(1)
    params = {k: v.clone() for k, v in tr.store.state_dict().items()}
    loss_ref, grads_ref, correct_ref = mlp_reference_step(params, xs[0].cpu(), ys[0].cpu())
    tr.step(xs[0], ys[0])
    torch.cuda.synchronize()
    assert abs(tr.read_loss() - loss_ref.item()) < 5e-2 * max(1.0, abs(loss_ref.item()))  # tf32 GEMM
    after = tr.store.state_dict()
    for name in tr.model.names():
        want = params[name] - 0.0005 * grads_ref[name]
        err = (after[name] - want).abs().max().item()
        assert err < 2e-4, (name, err)
    assert after["global_step"].item() == 1.0


@pytest.mark.parametrize("use_graph", [False, True])
def test_mlp_trains_and_graph_matches_eager(native_lib, use_graph):
    from b200ps.parallel.context import init_from_env
    from b200ps.train import MLPTrainer
    ctx = init_from_env(virtual_world=1)
    tr = MLPTrainer(ctx, batch_size=100, lr=0.01, mode="sync", use_graph=use_graph)
    xs, ys = _mnist_batches(50)
    first = None
    for i in range(150):
        tr.step(xs[i % 50], ys[i % 50])
        if i == 0:
            first = tr.read_loss()
    last = tr.read_loss()
    assert last < first and last == last
    acc, _ = tr.evaluate(xs.reshape(-1, 784), ys.reshape(-1))
    assert acc > 0.5, acc
    assert tr.global_step == 150


def test_resident_dataset_cursor_matches_copy_path(native_lib):
    """Walking a device-resident epoch through the in-graph cursor == feeding the same batches by copy."""
    from b200ps.parallel.context import init_from_env
    from b200ps.train import MLPTrainer
    xs, ys = _mnist_batches(7)
    a = MLPTrainer(init_from_env(virtual_world=1), batch_size=100, lr=0.01, mode="sync", use_graph=True)
    b = MLPTrainer(init_from_env(virtual_world=1), batch_size=100, lr=0.01, mode="sync", use_graph=True)
    a.attach_dataset(xs.reshape(-1, 784), ys.reshape(-1))
    a.run_resident(11)                        # one 8-step graph replay + 3 single steps
    for i in range(11, 17):
        a.step_resident()
    for i in range(17):                       # wraps around the 7-batch epoch twice
        b.step(xs[i % 7], ys[i % 7])
    torch.cuda.synchronize()
    assert a._cursor.item() == (17 * 100) % 700
    sa, sb = a.store.state_dict(), b.store.state_dict()
    for name in a.model.names():
        assert torch.equal(sa[name], sb[name]), name
    assert a.read_loss() == b.read_loss() and a.global_step == b.global_step == 17


def test_async_resident_cursor_matches_copy_path(native_lib):
    """Async mode: the cursor advanced by the push_async kernel walks the same batches as the copy path."""
    from b200ps.parallel.context import init_from_env
    from b200ps.train import MLPTrainer
    xs, ys = _mnist_batches(5)
    a = MLPTrainer(init_from_env(virtual_world=1), batch_size=100, lr=0.01, mode="async", use_graph=True)
    b = MLPTrainer(init_from_env(virtual_world=1), batch_size=100, lr=0.01, mode="async", use_graph=True)
    a.attach_dataset(xs.reshape(-1, 784), ys.reshape(-1))
    a.run_resident(13)
    for i in range(13):
        b.step(xs[i % 5], ys[i % 5])
    torch.cuda.synchronize()
    assert a._cursor.item() == (13 * 100) % 500
    sa, sb = a.store.state_dict(), b.store.state_dict()
    for name in a.model.names():
        assert torch.allclose(sa[name], sb[name], rtol=0, atol=1e-6), name
    assert a.global_step == b.global_step == 13


def test_native_host_step_matches_python_path(native_lib):
    """The one-call native e2e step (H2D + graph + D2H) gives the same losses as copy_/replay/item."""
    from b200ps.parallel.context import init_from_env
    from b200ps.train import MLPTrainer
    xs, ys = _mnist_batches(6)
    hx, hy = xs.cpu().pin_memory(), ys.cpu().pin_memory()
    a = MLPTrainer(init_from_env(virtual_world=1), batch_size=100, lr=0.01, mode="sync", use_graph=True)
    b = MLPTrainer(init_from_env(virtual_world=1), batch_size=100, lr=0.01, mode="sync", use_graph=True)
    la, lb, pending = [], [], []
    for i in range(12):
        pending.append(a.step_from_host_async(hx[i % 6], hy[i % 6]))     # native once the graph exists
        if len(pending) > 2:
            la.append(pending.pop(0).result())
        b.step(xs[i % 6], ys[i % 6])
        lb.append(b.read_loss())
    la += [h.result() for h in pending]
    assert a._feeder1 is not None and la == lb and a.global_step == b.global_step == 12


def test_chunked_host_feeder_and_fit_from_host(native_lib):
    """K steps per H2D copy / graph launch: same parameters and per-step losses as stepping one by one;
    tickets older than the ring are refused instead of returning another unit's losses."""
    import pytest as _pytest
    from b200ps.ops.native import NativeError
    from b200ps.parallel.context import init_from_env
    from b200ps.train import MLPTrainer
    n = 2 + 3 * 4 + 1                                   # warm-up + three chunks of 4 + a remainder batch
    xs, ys = _mnist_batches(n)
    hx, hy = xs.cpu().pin_memory(), ys.cpu().pin_memory()
    a = MLPTrainer(init_from_env(virtual_world=1), batch_size=100, lr=0.01, mode="sync", use_graph=True)
    b = MLPTrainer(init_from_env(virtual_world=1), batch_size=100, lr=0.01, mode="sync", use_graph=True)
    lb = []
    for i in range(n):
        b.step(xs[i], ys[i]); lb.append(b.read_loss())
    for i in range(2):
        a.step(xs[i], ys[i])
    la = []
    steps = a.fit_from_host(hx[2:].reshape(-1, 784), hy[2:].reshape(-1), steps_per_call=4,
                            on_loss=lambda i, v: la.append(v))
    torch.cuda.synchronize()
    assert steps == n - 2 and la == lb[2:] and a.global_step == b.global_step == n
    sa, sb = a.store.state_dict(), b.store.state_dict()
    for name in a.model.names():
        assert torch.equal(sa[name], sb[name]), name
    f = a.host_feeder(steps_per_call=1, slots=2)
    tickets = [f.submit(hx[i], hy[i]) for i in range(3)]
    assert tickets[2].result() == tickets[2].result()
    with _pytest.raises(NativeError):
        tickets[0].result()                              # slot reused by ticket 2
    f.close()


def test_sync_virtual_ranks_equal_big_batch(native_lib):
    """4 sync replicas x batch 100 == 1 worker x batch 400 (mean of means), up to fp32 rounding."""
    from b200ps.parallel.context import init_from_env
    from b200ps.train import MLPTrainer
    xs, ys = _mnist_batches(4)
    ctx4 = init_from_env(virtual_world=4)
    tr4 = MLPTrainer(ctx4, batch_size=100, lr=0.01, mode="sync", use_graph=False)
    for r in range(4):
        tr4.x[r].copy_(xs[r]); tr4.labels[r].copy_(ys[r])
    tr4.step()
    ctx1 = init_from_env(virtual_world=1)
    tr1 = MLPTrainer(ctx1, batch_size=400, lr=0.01, mode="sync", use_graph=False)
    tr1.step(xs.reshape(400, 784), ys.reshape(400))
    torch.cuda.synchronize()
    a, b = tr4.store.state_dict(), tr1.store.state_dict()
    for name in tr4.model.names():
        assert (a[name] - b[name]).abs().max().item() < 2e-4, name
        # every replica holds the same bytes after the push broadcast
        for r in range(1, 4):
            assert torch.equal(tr4.store.var(name, 0), tr4.store.var(name, r))
    assert tr4.global_step == 1


def test_async_virtual_ranks_count_every_update(native_lib):
    from b200ps.parallel.context import init_from_env
    from b200ps.train import MLPTrainer
    xs, ys = _mnist_batches(4)
    ctx = init_from_env(virtual_world=4)
    tr = MLPTrainer(ctx, batch_size=100, lr=0.01, mode="async", use_graph=False,
                    placement="round_robin_variable")
    for r in range(4):
        tr.x[r].copy_(xs[r]); tr.labels[r].copy_(ys[r])
    l0 = None
    for i in range(30):
        tr.step()
        if i == 0:
            l0 = tr.read_loss()
    assert tr.global_step == 4 * 30           # global_step counts ALL workers' updates (example.py:177-179)
    assert tr.read_loss() < l0
    # round-robin parity: creation order global_step, W1, W2, b1, b2 over 4 ps -> owners 1, 2, 3, 0
    m = tr.model
    assert [tr.store.owner_of(n) for n in (m.W1, m.W2, m.B1, m.B2)] == [1, 2, 3, 0]


def test_wide_mlp_small_matches_autograd(native_lib):
    from b200ps.parallel.context import init_from_env
    from b200ps.train import WideMLPTrainer
    from b200ps import ops
    B, L, D = 256, 3, 512
    for world in (1, 2):
        ctx = init_from_env(virtual_world=world)
        tr = WideMLPTrainer(ctx, batch_size=B, layers=L, hidden=D, lr=0.5, use_graph=False, fused_pull=True)
        xs = torch.empty(world, B, D, device=DEV, dtype=torch.bfloat16)
        ys = torch.empty(world, B, device=DEV, dtype=torch.int32)
        for r in range(world):
            ops.synth_dense_bf16(xs[r], ys[r], classes=D, seed=3, offset=r * B)
            tr.x[r].copy_(xs[r]); tr.labels[r].copy_(ys[r])
        before = tr.store.state_dict()
        # autograd reference on the bf16-rounded weights the kernels actually read
        ws = [before[tr.model.wname(l)].to(DEV).bfloat16().float().requires_grad_(True) for l in range(L)]
        bs = [before[tr.model.bname(l)].to(DEV).clone().requires_grad_(True) for l in range(L)]
        total = 0
        for r in range(world):
            a = xs[r].float()
            for l in range(L):
                z = a @ ws[l] + bs[l]
                a = torch.sigmoid(z).bfloat16().float() if l < L - 1 else z
            total = total + torch.nn.functional.cross_entropy(a, ys[r].long()) / world
        total.backward()
        tr.step()
        torch.cuda.synchronize()
        after = tr.store.state_dict()
        for l in range(L):
            want_w = before[tr.model.wname(l)] - 0.5 * ws[l].grad.cpu()
            want_b = before[tr.model.bname(l)] - 0.5 * bs[l].grad.cpu()
            ew = (after[tr.model.wname(l)] - want_w).abs().max().item()
            eb = (after[tr.model.bname(l)] - want_b).abs().max().item()
            scale_w = (0.5 * ws[l].grad).abs().max().item()
            assert ew < 0.05 * scale_w + 1e-5, (world, l, ew, scale_w)
            assert eb < 0.05 * (0.5 * bs[l].grad).abs().max().item() + 1e-5, (world, l, eb)
            # every rank's bf16 working copy agrees with the master on the rows it owns (the other
            # rows are re-pulled inside the next forward GEMM when the pull is fused)
            rows = D // world
            for r in range(world):
                w_dev = tr.store.var(tr.model.wname(l), r)[r * rows:(r + 1) * rows]
                assert torch.equal(w_dev.cpu(), after[tr.model.wname(l)][r * rows:(r + 1) * rows].bfloat16())
        if world > 1:
            assert tr.fused_pull
            tr.step()                                  # a second step consumes the fused pull of fresh weights
            torch.cuda.synchronize()
            assert tr.read_loss() == tr.read_loss() and tr.global_step == 2


def test_sparse_lr_trainer_learns(native_lib):
    from b200ps import ops
    from b200ps.metrics.auc import StreamingAUC
    from b200ps.parallel.context import init_from_env
    from b200ps.train import SparseLRTrainer
    F, B, nnz, spread = 1 << 12, 500, 40, 10
    # dedupe: the unique-id step (sparse_dedup.cu), a cooperative launch inside the captured step graph
    for mode, dedupe in (("async", False), ("sync", False), ("async", True), ("sync", True)):
        ctx = init_from_env(virtual_world=1)
        tr = SparseLRTrainer(ctx, features=F, lr=5.0, mode=mode, dedupe=dedupe)
        tr.store.sparse_shard(tr.model.W).mul_(0.01)       # start near zero so learning shows quickly
        batches = []
        for i in range(40):
            rp = torch.empty(B + 1, device=DEV, dtype=torch.int32)
            fid = torch.empty(B * (nnz + spread), device=DEV, dtype=torch.int64)
            fv = torch.empty(B * (nnz + spread), device=DEV); y = torch.empty(B, device=DEV)
            ops.synth_sparse(rp, fid, fv, y, nnz_mean=nnz, spread=spread, features=F, seed=5, offset=i * B)
            batches.append((rp, fid, fv, y))
        for epoch in range(6):
            for b in batches[:32]:
                tr.step(b)
        assert tr.global_step == 6 * 32
        auc = StreamingAUC(200, device=DEV)
        for b in batches[32:]:
            pred, _ = tr.predict(b)
            auc.update(pred, b[3])
        assert auc.result() > 0.65, (mode, dedupe, auc.result())
        if dedupe:
            distinct, nz = tr.model.last_unique_counts(batches[0][1].numel(), 0)
            assert 0 < distinct < nz                           # 2^12 features, ~20 k non-zeros: ids repeat


def test_sparse_wide_deep_trainer_learns_and_checkpoints(native_lib, tmp_path):
    """The wide & deep extension (lr2_debug.py:423-428) trains in both modes, on 1 and 2 (virtual) ranks,
    and its [F,D] table round-trips through the TF-bundle checkpoint."""
    from b200ps import ops
    from b200ps.io.saver import Saver
    from b200ps.metrics.auc import StreamingAUC
    from b200ps.parallel.context import init_from_env
    from b200ps.train import SparseWideDeepTrainer
    F, B, nnz, spread = 1 << 12, 500, 40, 10
    batches = []
    for i in range(40):
        rp = torch.empty(B + 1, device=DEV, dtype=torch.int32)
        fid = torch.empty(B * (nnz + spread), device=DEV, dtype=torch.int64)
        fv = torch.empty(B * (nnz + spread), device=DEV); y = torch.empty(B, device=DEV)
        ops.synth_sparse(rp, fid, fv, y, nnz_mean=nnz, spread=spread, features=F, seed=5, offset=i * B)
        batches.append((rp, fid, fv, y))
    # direct scatter of 512-byte rows, and the deferred form (scalar coefficients folded in by the shard owner)
    for mode, world, deferred in (("async", 1, False), ("sync", 1, True), ("async", 2, True)):
        ctx = init_from_env(virtual_world=world)
        tr = SparseWideDeepTrainer(ctx, features=F, dim=128, lr=2.0, mode=mode, deferred=deferred)
        assert tr.model.deferred == deferred and (("_pending/coefficients" in tr.store.sparse) == deferred)
        for epoch in range(6):
            for b in batches[:32]:
                tr.step(b)
        assert tr.global_step == 6 * 32
        auc = StreamingAUC(200, device=DEV)
        for b in batches[32:]:
            pred, _ = tr.predict(b)
            auc.update(pred, b[3])
        assert auc.result() > 0.65, (mode, world, auc.result())
    tr.flush()                                         # deferred updates folded in before the checkpoint
    state = tr.store.state_dict()
    assert state["deep/weights"].shape == (F, 128) and state["deep/weights_1"].shape == (128, 1)
    assert not any(k.startswith("_") for k in state)   # the pending tables are scratch, not model state
    from b200ps.io.saver import store_variables
    prefix = Saver().save(store_variables(tr.store), str(tmp_path / "model.ckpt"), global_step=tr.global_step)
    back = Saver().restore(prefix)
    tr2 = SparseWideDeepTrainer(init_from_env(virtual_world=2), features=F, dim=128, lr=2.0)
    tr2.store.load_state_dict({k: torch.as_tensor(v) for k, v in back.items()})
    p1, _ = tr.predict(batches[33]); p2, _ = tr2.predict(batches[33])
    assert torch.equal(p1, p2)


def test_native_loader_feeds_sparse_trainer(native_lib, tmp_path):
    """files -> native reader thread -> packed pinned batches -> HostFeeder (one H2D copy + one graph per step):
    same losses and weights as formatting the same lines in Python and stepping batch by batch."""
    import numpy as np
    from b200ps.data.libsvm import Sample
    from b200ps.data.native_loader import NativeLibsvmLoader
    from b200ps.data.synthetic import write_libsvm_files
    from b200ps.parallel.context import init_from_env
    from b200ps.train import SparseLRTrainer
    F, B = 5000, 100
    files = write_libsvm_files(str(tmp_path / "train"), 2, 450, nnz_mean=20, spread=5, features=F, seed=3)
    # sync mode: forward on frozen weights, then the update — reproducible up to the order of float atomics
    a = SparseLRTrainer(init_from_env(virtual_world=1), features=F, lr=0.5, mode="sync")
    b = SparseLRTrainer(init_from_env(virtual_world=1), features=F, lr=0.5, mode="sync", use_graph=False)
    loader = NativeLibsvmLoader(files, batch_size=B, threads=1, max_nnz_per_row=32, epochs=2)
    assert loader.pinned
    la = []
    steps = a.fit_from_loader(loader, on_loss=lambda i, v: la.append(v))
    assert steps == 2 * (900 // B) == loader.stats()["batches"] and a.global_step == steps
    samples = [Sample().parse_line_libsvm(l) for f in files for l in open(f)]
    lb = []
    for _ in range(2):
        for lo in range(0, 900 - B + 1, B):
            batch = tuple(torch.from_numpy(x).to(DEV) for x in Sample.format_samples_csr(samples[lo:lo + B]))
            b.step(batch)
            lb.append(b.read_loss())
    np.testing.assert_allclose(la, lb, rtol=1e-4, atol=1e-5)
    wa, wb = a.store.state_dict()[a.model.W], b.store.state_dict()[b.model.W]
    assert torch.allclose(wa, wb, rtol=1e-4, atol=1e-5)
    loader.close()


def test_cluster_step_trainer_matches_default_path(native_lib, monkeypatch):
    """The opt-in resident-cluster step (K14, B200PS_CLUSTER_STEP=1) drives the same trainer APIs — single steps,
    multi-step resident launches, chunked host feeding — and lands within fp32 rounding of the default K13 path."""
    from b200ps import ops
    from b200ps.parallel.context import init_from_env
    from b200ps.train import MLPTrainer
    if not ops.mlp_cluster_supported(100, 784, 100, 10):
        pytest.skip("16-CTA clusters are not schedulable on this device")
    xs, ys = _mnist_batches(7)
    a = MLPTrainer(init_from_env(virtual_world=1), batch_size=100, lr=0.01, mode="sync", use_graph=True)
    monkeypatch.setenv("B200PS_CLUSTER_STEP", "1")
    b = MLPTrainer(init_from_env(virtual_world=1), batch_size=100, lr=0.01, mode="sync", use_graph=True)
    assert b.cluster and not a.cluster
    for tr in (a, b):
        tr.attach_dataset(xs.reshape(-1, 784), ys.reshape(-1))
        for i in range(3):
            tr.step(xs[i], ys[i])
        tr.run_resident(11)
        tr.step_resident()
    torch.cuda.synchronize()
    assert a.global_step == b.global_step == 15 and a._cursor.item() == b._cursor.item()
    sa, sb = a.store.state_dict(), b.store.state_dict()
    for name in a.model.names():
        assert torch.allclose(sa[name], sb[name], atol=2e-5, rtol=1e-4), name
    assert abs(a.read_loss() - b.read_loss()) < 1e-4
    hx, hy = xs.reshape(-1, 784)[:400].cpu().pin_memory(), ys.reshape(-1)[:400].cpu().pin_memory()
    la, lb = [], []
    a.fit_from_host(hx, hy, steps_per_call=4, on_loss=lambda i, v: la.append(v))
    b.fit_from_host(hx, hy, steps_per_call=4, on_loss=lambda i, v: lb.append(v))
    assert len(la) == len(lb) == 4 and max(abs(p - q) for p, q in zip(la, lb)) < 1e-4
