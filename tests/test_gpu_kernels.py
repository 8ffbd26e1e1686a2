"""Non-GEMM native kernels vs PyTorch fp32/fp64 references (K1, K3-K6, K9-K12)."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def test_mlp_tail_matches_autograd(native_lib):
    from b200ps import ops
    g = torch.Generator(device=DEV).manual_seed(0)
    B, H, Cn = 100, 100, 10
    z2 = torch.randn(B, H, device=DEV, generator=g)
    b1 = torch.randn(H, device=DEV, generator=g) * 0.1
    w2 = torch.randn(H, Cn, device=DEV, generator=g) * 0.3
    b2 = torch.randn(Cn, device=DEV, generator=g) * 0.1
    labels = torch.randint(0, Cn, (B,), device=DEV, generator=g, dtype=torch.int32)

    z2r = z2.clone().requires_grad_(True)
    b1r, w2r, b2r = (t.clone().requires_grad_(True) for t in (b1, w2, b2))
    a2 = torch.sigmoid(z2r + b1r)
    logits = a2 @ w2r + b2r
    loss = torch.nn.functional.cross_entropy(logits, labels.long())
    loss.backward()

    loss_sum = torch.zeros(1, device=DEV)
    correct = torch.zeros(1, device=DEV, dtype=torch.int32)
    dz2 = torch.empty(B, H, device=DEV)
    dw2 = torch.empty(H, Cn, device=DEV)
    db2 = torch.empty(Cn, device=DEV)
    db1 = torch.empty(H, device=DEV)
    ops.mlp_tail(z2, b1, w2, b2, labels, loss_sum=loss_sum, correct=correct, dz2=dz2, dw2=dw2,
                 db2=db2, db1=db1)
    torch.cuda.synchronize()
    assert abs(loss_sum.item() - loss.item()) < 1e-4
    assert correct.item() == (logits.argmax(1) == labels).sum().item()
    assert torch.allclose(dz2, z2r.grad, atol=1e-6, rtol=1e-4)
    assert torch.allclose(dw2, w2r.grad, atol=1e-6, rtol=1e-4)
    assert torch.allclose(db2, b2r.grad, atol=1e-6, rtol=1e-4)
    assert torch.allclose(db1, b1r.grad, atol=1e-6, rtol=1e-4)


@pytest.mark.parametrize("B,D,H,Cn", [(100, 784, 100, 10), (37, 48, 20, 3), (128, 64, 128, 16), (8, 16, 4, 2)])
def test_mlp_fused_step_matches_fp32_reference(native_lib, B, D, H, Cn):
    """K13, both forms: gradients for a push, and the in-place single-replica update with global_step and
    the resident-dataset cursor.  Exact-fp32 kernel vs the float64 restatement of the same step."""
    from b200ps import ops
    g = torch.Generator(device=DEV).manual_seed(4)
    rows = 3 * B
    x = torch.rand(rows, D, device=DEV, generator=g)
    labels = torch.randint(0, Cn, (rows,), device=DEV, generator=g, dtype=torch.int32)
    w1 = torch.randn(D, H, device=DEV, generator=g) * 0.2
    b1 = torch.randn(H, device=DEV, generator=g) * 0.1
    w2 = torch.randn(H, Cn, device=DEV, generator=g) * 0.5
    b2 = torch.randn(Cn, device=DEV, generator=g) * 0.1

    def ref(xb, yb):
        W1, B1, W2, B2 = (t.double().clone().requires_grad_(True) for t in (w1, b1, w2, b2))
        logits = torch.sigmoid(xb.double() @ W1 + B1) @ W2 + B2
        loss = torch.nn.functional.cross_entropy(logits, yb.long())
        loss.backward()
        return loss.item(), int((logits.argmax(1) == yb).sum()), (W1.grad, B1.grad, W2.grad, B2.grad)

    cursor = torch.tensor([B], device=DEV, dtype=torch.int32)           # second batch of the resident rows
    loss = torch.zeros(1, device=DEV); correct = torch.zeros(1, device=DEV, dtype=torch.int32)
    grads = (torch.empty_like(w1), torch.empty_like(b1), torch.empty_like(w2), torch.empty_like(b2))
    before = [t.clone() for t in (w1, b1, w2, b2)]
    ops.mlp_fused_step(x, labels, w1, b1, w2, b2, loss=loss, correct=correct, grads=grads, batch=B, cursor=cursor)
    torch.cuda.synchronize()
    l_ref, c_ref, g_ref = ref(x[B:2 * B], labels[B:2 * B])
    assert abs(loss.item() - l_ref) < 2e-5 * max(1.0, abs(l_ref)) and correct.item() == c_ref
    for got, want in zip(grads, g_ref):
        assert torch.allclose(got.double(), want, atol=2e-6, rtol=2e-4)
    for t, b in zip((w1, b1, w2, b2), before):
        assert torch.equal(t, b)                                        # parameters untouched, cursor too
    assert cursor.item() == B
    # in-place form: two consecutive steps walk the cursor and count global_step
    gstep = torch.zeros(1, device=DEV, dtype=torch.int64)
    lr = 0.1
    want = [t.double() - lr * gr for t, gr in zip(before, g_ref)]
    for _ in range(2):
        ops.mlp_fused_step(x, labels, w1, b1, w2, b2, loss=loss, correct=correct, batch=B, cursor=cursor, lr=lr,
                           global_step=gstep, cursor_rw=cursor, cursor_step=B, cursor_wrap=3 * B)
        if _ == 0:
            torch.cuda.synchronize()
            for t, w in zip((w1, b1, w2, b2), want):
                assert torch.allclose(t.double(), w, atol=2e-6, rtol=1e-5)
    torch.cuda.synchronize()
    assert gstep.item() == 2 and cursor.item() == 0                     # B -> 2B -> wraps to 0
    # bit-reproducible: the same two steps from the same start give the same bytes
    again = [b.clone() for b in before]
    cursor.fill_(B)
    for _ in range(2):
        ops.mlp_fused_step(x, labels, *again, loss=loss, correct=correct, batch=B, cursor=cursor, lr=lr,
                           global_step=gstep, cursor_rw=cursor, cursor_step=B, cursor_wrap=3 * B)
    torch.cuda.synchronize()
    for t, a in zip((w1, b1, w2, b2), again):
        assert torch.equal(t, a)


def test_mlp_tail_eval_large_batch(native_lib):
    from b200ps import ops
    g = torch.Generator(device=DEV).manual_seed(1)
    B, H, Cn = 10000, 100, 10
    z2 = torch.randn(B, H, device=DEV, generator=g)
    b1 = torch.zeros(H, device=DEV)
    w2 = torch.randn(H, Cn, device=DEV, generator=g)
    b2 = torch.zeros(Cn, device=DEV)
    labels = torch.randint(0, Cn, (B,), device=DEV, generator=g, dtype=torch.int32)
    loss_sum = torch.zeros(1, device=DEV)
    correct = torch.zeros(1, device=DEV, dtype=torch.int32)
    ops.mlp_tail(z2, b1, w2, b2, labels, loss_sum=loss_sum, correct=correct, train=False)
    logits = torch.sigmoid(z2) @ w2
    ref = torch.nn.functional.cross_entropy(logits, labels.long())
    assert abs(loss_sum.item() - ref.item()) < 1e-3
    assert correct.item() == (logits.argmax(1) == labels).sum().item()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_softmax_ce(native_lib, dtype):
    from b200ps import ops
    g = torch.Generator(device=DEV).manual_seed(2)
    B, Cn = 64, 8192
    logits = (torch.randn(B, Cn, device=DEV, generator=g) * 2).to(dtype)
    labels = torch.randint(0, Cn, (B,), device=DEV, generator=g, dtype=torch.int32)
    lr = logits.float().clone().requires_grad_(True)
    loss = torch.nn.functional.cross_entropy(lr, labels.long())
    loss.backward()
    d = torch.empty_like(logits)
    loss_sum = torch.zeros(1, device=DEV)
    correct = torch.zeros(1, device=DEV, dtype=torch.int32)
    ops.softmax_ce(logits, labels, dlogits=d, loss_sum=loss_sum, correct=correct)
    assert abs(loss_sum.item() - loss.item()) < 2e-3
    assert correct.item() == (logits.float().argmax(1) == labels).sum().item()
    tol = 1e-6 if dtype == torch.float32 else 1e-4
    assert (d.float() - lr.grad).abs().max().item() < tol


def test_colsum_sgd_adam(native_lib):
    from b200ps import ops
    g = torch.Generator(device=DEV).manual_seed(3)
    x = torch.randn(1000, 300, device=DEV, generator=g).bfloat16()
    out = torch.empty(300, device=DEV)
    ops.colsum(x, out, scale=0.5)
    assert torch.allclose(out, 0.5 * x.float().sum(0), atol=1e-2, rtol=1e-3)

    p = torch.randn(10007, device=DEV, generator=g)
    gr = torch.randn(10007, device=DEV, generator=g)
    want = p - 0.05 * gr
    lowp = torch.empty(10007, device=DEV, dtype=torch.bfloat16)
    ops.sgd_apply(p, gr, 0.05, lowp=lowp)
    assert torch.allclose(p, want, atol=1e-7)
    assert torch.equal(lowp, p.bfloat16())

    # Adam vs torch.optim.Adam (same formula as TF's ApplyAdam up to eps placement: use TF's)
    w = torch.randn(513, device=DEV, generator=g)
    m = torch.zeros_like(w); v = torch.zeros_like(w)
    bp = torch.tensor([0.9, 0.999], device=DEV)
    wr = w.double().clone(); mr = torch.zeros_like(wr); vr = torch.zeros_like(wr)
    for t in range(1, 4):
        gr = torch.randn(513, device=DEV, generator=g)
        ops.adam_apply(w, gr, m, v, bp, lr=0.01)
        ops.adam_advance(bp)
        gd = gr.double()
        mr = 0.9 * mr + 0.1 * gd
        vr = 0.999 * vr + 0.001 * gd * gd
        lr_t = 0.01 * math.sqrt(1 - 0.999 ** t) / (1 - 0.9 ** t)
        wr = wr - lr_t * mr / (vr.sqrt() + 1e-8)
    assert torch.allclose(w.double(), wr, atol=1e-5)


def test_linreg(native_lib):
    from b200ps import ops
    g = torch.Generator(device=DEV).manual_seed(4)
    x = torch.rand(100, 1, device=DEV, generator=g) * 100
    y = x + 20 * torch.sin(x / 10)
    w = torch.randn(1, 1, device=DEV, generator=g); b = torch.zeros(1, device=DEV)
    wr = w.clone().requires_grad_(True); br = b.clone().requires_grad_(True)
    loss = (((y - (x @ wr + br)) ** 2) / 1000).sum()
    loss.backward()
    lo = torch.zeros(1, device=DEV); gw = torch.zeros(1, device=DEV); gb = torch.zeros(1, device=DEV)
    ops.linreg_fwd_bwd(x, y.reshape(-1), w, b, loss=lo, gw=gw, gb=gb, inv_n=1 / 1000)
    assert torch.allclose(lo, loss.detach().reshape(1), rtol=1e-4)
    assert torch.allclose(gw, wr.grad.reshape(1), rtol=1e-4)
    assert torch.allclose(gb, br.grad.reshape(1), rtol=1e-4)


def test_synth_generators(native_lib):
    from b200ps import ops
    x = torch.empty(512, 784, device=DEV); lab = torch.empty(512, device=DEV, dtype=torch.int32)
    oh = torch.empty(512, 10, device=DEV)
    ops.synth_mnist(x, lab, onehot=oh, seed=7, offset=100)
    assert 0 <= x.min().item() and x.max().item() <= 1
    assert torch.equal(oh.argmax(1).int(), lab) and set(lab.tolist()) <= set(range(10))
    x2 = torch.empty(256, 784, device=DEV); lab2 = torch.empty(256, device=DEV, dtype=torch.int32)
    ops.synth_mnist(x2, lab2, seed=7, offset=356)      # rows 256.. of the first call
    assert torch.equal(x2, x[256:]) and torch.equal(lab2, lab[256:])

    xb = torch.empty(64, 256, device=DEV, dtype=torch.bfloat16)
    lb = torch.empty(64, device=DEV, dtype=torch.int32)
    ops.synth_dense_bf16(xb, lb, classes=256, seed=3)
    assert xb.float().abs().max().item() <= 1.0 + 1e-3 and lb.max().item() < 256

    B, nnz, spread, F = 500, 100, 20, 1 << 24
    rp = torch.empty(B + 1, device=DEV, dtype=torch.int32)
    fid = torch.empty(B * (nnz + spread), device=DEV, dtype=torch.int64)
    fv = torch.empty(B * (nnz + spread), device=DEV)
    y = torch.empty(B, device=DEV)
    ops.synth_sparse(rp, fid, fv, y, nnz_mean=nnz, spread=spread, features=F, seed=5)
    lens = (rp[1:] - rp[:-1])
    assert rp[0].item() == 0 and lens.min().item() >= nnz - spread and lens.max().item() <= nnz + spread
    n = rp[-1].item()
    assert 0 <= fid[:n].min().item() and fid[:n].max().item() < F
    assert set(y.tolist()) <= {0.0, 1.0} and 0.2 < y.mean().item() < 0.8

    w = torch.empty(100003, device=DEV)
    ops.init_normal(w, seed=1, stream_id=3)
    assert abs(w.mean().item()) < 0.02 and abs(w.std().item() - 1) < 0.02


def _sparse_batch(B=500, nnz=100, spread=20, F=1 << 20, seed=5):
    from b200ps import ops
    rp = torch.empty(B + 1, device=DEV, dtype=torch.int32)
    fid = torch.empty(B * (nnz + spread), device=DEV, dtype=torch.int64)
    fv = torch.empty(B * (nnz + spread), device=DEV)
    y = torch.empty(B, device=DEV)
    ops.synth_sparse(rp, fid, fv, y, nnz_mean=nnz, spread=spread, features=F, seed=seed)
    return rp, fid, fv, y


@pytest.mark.parametrize("world", [1, 4])
def test_sparse_lr_vs_dense_reference(native_lib, world):
    from b200ps import ops
    F = 1 << 20
    rp, fid, fv, y = _sparse_batch(F=F)
    B = y.numel(); n = rp[-1].item()
    g = torch.Generator(device=DEV).manual_seed(9)
    W = torch.randn(F, device=DEV, generator=g) * 0.1
    rows = F // world
    shards = [W[r * rows:(r + 1) * rows].clone() for r in range(world)]
    table = ops.device_ptr_table([s.data_ptr() for s in shards], DEV)
    bias = torch.full((1,), 0.25, device=DEV)
    # float64 reference
    row_of = torch.repeat_interleave(torch.arange(B, device=DEV), (rp[1:] - rp[:-1]).long())
    z = torch.zeros(B, device=DEV, dtype=torch.float64).index_add_(
        0, row_of, W.double()[fid[:n]] * fv[:n].double()) + 0.25
    p_ref = torch.sigmoid(z)
    loss_ref = torch.nn.functional.binary_cross_entropy_with_logits(z, y.double())
    dz_ref = (p_ref - y.double()) / B
    lr = 0.5
    W_ref = W.double().index_add(0, fid[:n], -lr * dz_ref[row_of] * fv[:n].double())
    b_ref = 0.25 - lr * dz_ref.sum().item()

    # forward only: exact predictions / loss / dz on frozen weights
    pred = torch.empty(B, device=DEV); loss = torch.zeros(1, device=DEV); dz = torch.empty(B, device=DEV)
    gstep = torch.zeros(1, device=DEV, dtype=torch.int64)
    ops.sparse_lr(rp, fid, fv, y, table, bias, rows_per_shard=rows, pred=pred, dz_out=dz, loss_sum=loss,
                  forward=True, update=False)
    torch.cuda.synchronize()
    assert torch.allclose(pred.double(), p_ref, atol=1e-5)
    assert abs(loss.item() - loss_ref.item()) < 1e-5
    assert torch.allclose(dz.double(), dz_ref, atol=1e-8)
    assert torch.equal(torch.cat(shards), W) and gstep.item() == 0
    # update only (sync-mode split): the scatter-add must conserve every contribution
    ops.sparse_lr(rp, fid, fv, y, table, bias, rows_per_shard=rows, dz_in=dz, global_step=gstep,
                  forward=False, update=True, lr=lr)
    torch.cuda.synchronize()
    got = torch.cat(shards).double()
    assert (got - W_ref).abs().max().item() < 1e-5
    assert abs(bias.item() - b_ref) < 1e-5 and gstep.item() == 1
    # fused Hogwild step: later warps may read rows earlier warps already updated, so only
    # closeness (not equality) to the frozen-weight result is expected
    shards2 = [W[r * rows:(r + 1) * rows].clone() for r in range(world)]
    table2 = ops.device_ptr_table([s.data_ptr() for s in shards2], DEV)
    bias2 = torch.full((1,), 0.25, device=DEV)
    ops.sparse_lr(rp, fid, fv, y, table2, bias2, rows_per_shard=rows, loss_sum=loss, global_step=gstep,
                  forward=True, update=True, lr=lr)
    torch.cuda.synchronize()
    assert (torch.cat(shards2).double() - W_ref).abs().max().item() < 1e-3 and gstep.item() == 2


def _skewed_batch(F, seed, B=500, nnz=100, spread=20):
    """A click-log-like batch: ids drawn as F * u^6 (a few hot ids in most rows), plus two ids outside the range."""
    rp, fid, fv, y = _sparse_batch(B=B, nnz=nnz, spread=spread, F=F, seed=seed)
    n = rp[-1].item()
    g = torch.Generator(device=DEV).manual_seed(seed)
    u = torch.rand(n, device=DEV, generator=g, dtype=torch.float64)
    fid[:n] = (u.pow(6) * F).long().clamp_(0, F - 1)
    return rp, fid, fv, y


@pytest.mark.parametrize("world", [1, 4])
def test_sparse_lr_unique_ids_vs_dense_reference(native_lib, world):
    """K4u/K5u (sparse_dedup.cu): the unique-id step against a float64 dense reference on batches full of repeated
    ids — forward, the synchronous split, the fused step (exact here: every weight is read before any is written),
    the distinct-id count, refused ids, and a second and third batch through the same never-cleared table."""
    from b200ps import ops
    from b200ps.models.sparse_lr import SparseLR
    F = 1 << 20
    rows = F // world
    g = torch.Generator(device=DEV).manual_seed(9)
    W = torch.randn(F, device=DEV, generator=g) * 0.1
    shards = [W[r * rows:(r + 1) * rows].clone() for r in range(world)]
    table = ops.device_ptr_table([s.data_ptr() for s in shards], DEV)
    bias = torch.full((1,), 0.25, device=DEV)
    gstep = torch.zeros(1, device=DEV, dtype=torch.int64)
    model = SparseLR(F, dedupe=True)
    lr = 0.5
    W_ref, b_ref = W.double().clone(), 0.25
    ops.sparse_bad_ids(reset=True)
    for it, mode in enumerate(["split", "fused", "fused"]):
        rp, fid, fv, y = _skewed_batch(F, seed=11 + it)
        B = y.numel(); n = rp[-1].item()
        if it == 1:
            fid[3] = F + 5; fid[n - 1] = -2                  # refused: contribute nothing, counted
        ok = (fid[:n] >= 0) & (fid[:n] < F)
        sc = model._dedup_scratch(fid.numel(), DEV, 0)
        row_of = torch.repeat_interleave(torch.arange(B, device=DEV), (rp[1:] - rp[:-1]).long())
        fsafe = fid[:n].clamp(0, F - 1)
        contrib = torch.where(ok, W_ref[fsafe] * fv[:n].double(), torch.zeros((), device=DEV, dtype=torch.float64))
        z = torch.zeros(B, device=DEV, dtype=torch.float64).index_add_(0, row_of, contrib) + b_ref
        p_ref = torch.sigmoid(z)
        loss_ref = torch.nn.functional.binary_cross_entropy_with_logits(z, y.double())
        dz_ref = (p_ref - y.double()) / B
        upd = torch.where(ok, -lr * dz_ref[row_of] * fv[:n].double(), torch.zeros((), device=DEV, dtype=torch.float64))
        W_ref = W_ref.index_add(0, fsafe, upd)
        b_ref = b_ref - lr * dz_ref.sum().item()
        pred = torch.empty(B, device=DEV); loss = torch.zeros(1, device=DEV); dz = torch.empty(B, device=DEV)
        kw = dict(scratch=sc, rows_per_shard=rows, features=F)
        if mode == "split":
            before = torch.cat(shards).clone()
            ops.sparse_lr_unique(rp, fid, fv, y, table, bias, pred=pred, dz_out=dz, loss_sum=loss, forward=True,
                                 update=False, finish=False, **kw)
            torch.cuda.synchronize()
            assert torch.allclose(pred.double(), p_ref, atol=1e-5) and abs(loss.item() - loss_ref.item()) < 1e-5
            assert torch.allclose(dz.double(), dz_ref, atol=1e-8) and torch.equal(torch.cat(shards), before)
            ops.sparse_lr_unique(rp, fid, fv, y, table, bias, dz_in=dz, global_step=gstep, forward=False, update=True,
                                 lr=lr, unique=False, **kw)
        else:
            ops.sparse_lr_unique(rp, fid, fv, y, table, bias, pred=pred, loss_sum=loss, global_step=gstep,
                                 forward=True, update=True, lr=lr, **kw)
            torch.cuda.synchronize()
            assert torch.allclose(pred.double(), p_ref, atol=1e-5) and abs(loss.item() - loss_ref.item()) < 1e-5
        torch.cuda.synchronize()
        got = torch.cat(shards).double()
        assert (got - W_ref).abs().max().item() < 2e-5, (it, (got - W_ref).abs().max().item())
        assert abs(bias.item() - b_ref) < 1e-5 and gstep.item() == it + 1
        distinct, nz = model.last_unique_counts(fid.numel(), 0)
        assert distinct == torch.unique(fid[:n][ok]).numel() and nz == n
        assert distinct < 0.8 * n                              # the batch really repeats ids
        W_ref = got                                            # carry the fp32 state forward
    assert ops.sparse_bad_ids(reset=True) == 2


@pytest.mark.parametrize("world,dim,wide", [(1, 128, True), (4, 128, True), (2, 256, False), (2, 512, True)])
def test_sparse_wide_deep_vs_float64_reference(native_lib, world, dim, wide):
    """K4w/K5w: forward exact on frozen tables; update-only conserves every contribution of all four
    gradients; the fused Hogwild launch lands close to the frozen-weight result."""
    from b200ps import ops
    from b200ps.models.wide_deep import wide_deep_reference
    F, B = 1 << 14, 300
    rp, fid, fv, y = _sparse_batch(B=B, nnz=40, spread=10, F=F)
    g = torch.Generator(device=DEV).manual_seed(9)
    W = torch.randn(F, device=DEV, generator=g) * 0.1
    E = torch.randn(F, dim, device=DEV, generator=g) * 0.1
    w2 = torch.randn(dim, device=DEV, generator=g) * 0.3
    rows, rows2 = F // world, dim // world

    def tables():
        ws = [W[r * rows:(r + 1) * rows].clone() for r in range(world)]
        es = [E[r * rows:(r + 1) * rows].clone() for r in range(world)]
        w2s = [w2[r * rows2:(r + 1) * rows2].clone() for r in range(world)]
        kw = dict(wide_table=ops.device_ptr_table([t.data_ptr() for t in ws], DEV) if wide else None,
                  wide_rows_per_shard=rows if wide else 0,
                  emb_table=ops.device_ptr_table([t.data_ptr() for t in es], DEV), emb_rows_per_shard=rows, dim=dim,
                  w2_table=ops.device_ptr_table([t.data_ptr() for t in w2s], DEV), w2_rows_per_shard=rows2,
                  bias=torch.full((1,), 0.25, device=DEV))
        return ws, es, w2s, kw

    lr = 0.5
    loss_ref, p_ref, gr = wide_deep_reference(W.cpu() if wide else None, E.cpu(), w2.cpu(), torch.tensor(0.25),
                                              rp.cpu(), fid.cpu(), fv.cpu(), y.cpu())
    ws, es, w2s, kw = tables()
    pred = torch.empty(B, device=DEV); loss = torch.zeros(1, device=DEV); dz = torch.empty(B, device=DEV)
    gstep = torch.zeros(1, device=DEV, dtype=torch.int64)
    ops.sparse_wd(rp, fid, fv, y, **kw, pred=pred, dz_out=dz, loss_sum=loss, forward=True, update=False)
    torch.cuda.synchronize()
    assert torch.allclose(pred.cpu().double(), p_ref, atol=2e-5)
    assert abs(loss.item() - loss_ref.item()) < 2e-5
    assert torch.equal(torch.cat(es), E) and torch.equal(torch.cat(w2s), w2)
    ops.sparse_wd(rp, fid, fv, y, **kw, dz_in=dz, global_step=gstep, forward=False, update=True, lr=lr)
    torch.cuda.synchronize()
    assert (torch.cat(es).cpu().double() - (E.cpu().double() - lr * gr["E"])).abs().max().item() < 1e-5
    # (the w2 gradient is a sum of 300 signed dz*h terms accumulated in fp32 across warps and CTAs)
    assert (torch.cat(w2s).cpu().double() - (w2.cpu().double() - lr * gr["w2"])).abs().max().item() < 5e-5
    if wide:
        assert (torch.cat(ws).cpu().double() - (W.cpu().double() - lr * gr["W"])).abs().max().item() < 1e-5
    assert abs(kw["bias"].item() - (0.25 - lr * gr["b"].item())) < 1e-5 and gstep.item() == 1
    ws, es, w2s, kw = tables()
    ops.sparse_wd(rp, fid, fv, y, **kw, loss_sum=loss, global_step=gstep, forward=True, update=True, lr=lr)
    torch.cuda.synchronize()
    assert (torch.cat(es).cpu().double() - (E.cpu().double() - lr * gr["E"])).abs().max().item() < 1e-3
    assert (torch.cat(w2s).cpu().double() - (w2.cpu().double() - lr * gr["w2"])).abs().max().item() < 1e-3
    assert gstep.item() == 2
    # deferred mode: workers add one scalar per non-zero into the pending table, the owners fold it in
    # (rows first, then w2): same result as the direct scatter, and the tables are untouched until the fold
    ws, es, w2s, kw = tables()
    pend = [torch.zeros(rows, device=DEV) for _ in range(world)]
    pend2 = [torch.zeros(rows2, device=DEV) for _ in range(world)]
    dk = dict(pend_table=ops.device_ptr_table([t.data_ptr() for t in pend], DEV), pend_rows_per_shard=rows,
              pend2_table=ops.device_ptr_table([t.data_ptr() for t in pend2], DEV), pend2_rows_per_shard=rows2)
    ops.sparse_wd(rp, fid, fv, y, **kw, **dk, dz_in=dz, global_step=gstep, forward=False, update=True, lr=lr)
    torch.cuda.synchronize()
    assert torch.equal(torch.cat(es), E) and torch.equal(torch.cat(w2s), w2) and torch.cat(pend).abs().sum() > 0
    for phase in (1, 2):
        for r in range(world):
            ops.sparse_wd_apply(pend[r], ws[r] if wide else None, es[r], rows, dim, kw["w2_table"], rows2, pend2[r], w2s[r],
                                rows2, phase)
    torch.cuda.synchronize()
    assert (torch.cat(es).cpu().double() - (E.cpu().double() - lr * gr["E"])).abs().max().item() < 1e-5
    assert (torch.cat(w2s).cpu().double() - (w2.cpu().double() - lr * gr["w2"])).abs().max().item() < 5e-5
    if wide:
        assert (torch.cat(ws).cpu().double() - (W.cpu().double() - lr * gr["W"])).abs().max().item() < 1e-5
    assert torch.cat(pend).abs().sum().item() == 0 and torch.cat(pend2).abs().sum().item() == 0


def test_auc_hist(native_lib):
    from b200ps import ops
    from b200ps.metrics.auc import StreamingAUC, auc_from_counts_reference
    g = torch.Generator(device=DEV).manual_seed(11)
    pred = torch.rand(5000, device=DEV, generator=g)
    pred[:7] = torch.tensor([0., 1., 0.5, 1 / 199, 2 / 199, 198 / 199, 0.25], device=DEV)
    y = (torch.rand(5000, device=DEV, generator=g) < pred).float()
    m = StreamingAUC(num_thresholds=200, device=DEV)
    m.update(pred[:3000], y[:3000]); m.update(pred[3000:], y[3000:])
    tp, fn, tn, fp = m.counts()
    rtp, rfn, rtn, rfp = auc_from_counts_reference(pred.cpu().numpy(), y.cpu().numpy(), 200)
    assert (tp == rtp).all() and (fn == rfn).all() and (tn == rtn).all() and (fp == rfp).all()
    assert 0.6 < m.result() < 1.0


def _virtual_ranks(world, n, dtype):
    """One GPU, `world` emulated ranks: full-size working copies, grads, pads on the same device."""
    from b200ps import ops
    g = torch.Generator(device=DEV).manual_seed(13)
    shard = ((n + world - 1) // world + 7) // 8 * 8
    total = shard * world
    bounds = torch.tensor([min(r * shard, total) for r in range(world + 1)], device=DEV, dtype=torch.int64)
    init = torch.randn(total, device=DEV, generator=g)
    work = [init.to(dtype).clone() for _ in range(world)]
    if dtype == torch.float32:
        masters = work                                     # master lives in place in the working copy
        master_ptrs = [w.data_ptr() for w in work]
    else:
        masters = [init[r * shard:(r + 1) * shard].clone() for r in range(world)]
        master_ptrs = [m.data_ptr() - 4 * r * shard for r, m in enumerate(masters)]
    grads = [(torch.randn(total, device=DEV, generator=g)).to(dtype) for _ in range(world)]
    pads = [torch.zeros(ops.signal_pad_words(), device=DEV, dtype=torch.int32) for _ in range(world)]
    gsteps = [torch.zeros(1, device=DEV, dtype=torch.int64) for _ in range(world)]
    T = lambda ts: ops.device_ptr_table([t if isinstance(t, int) else t.data_ptr() for t in ts], DEV)
    return dict(shard=shard, total=total, bounds=bounds, init=init, work=work, masters=masters,
                grads=grads, pads=pads, gsteps=gsteps, t_pads=T(pads), t_masters=T(master_ptrs),
                t_outs=T(work), t_grads=T(grads), t_gsteps=T(gsteps))


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("world,R", [(1, 1), (4, 4), (4, 3)])
def test_push_sync_virtual_ranks(native_lib, dtype, world, R):
    from b200ps import ops
    v = _virtual_ranks(world, 79510, dtype)
    lr = 0.1
    expected = v["init"].clone()
    for step in range(1, 3):
        ops.push_sync_sgd(v["t_pads"], v["t_masters"], v["t_outs"], v["t_grads"], v["bounds"], v["t_gsteps"],
                          world=world, rank_base=0, local_ranks=world, max_shard_elems=v["shard"],
                          dtype=dtype, replicas_to_aggregate=R, lr=lr)
        torch.cuda.synchronize()
        sel = [r for r in range(world) if R >= world or ((r - step % world) % world) < R]
        s = torch.zeros_like(expected)
        for r in sel:                                   # same fixed order as the kernel
            s += v["grads"][r].float()
        expected = expected - (lr / min(R, world)) * s
        for r in range(world):
            assert torch.equal(v["work"][r], v["work"][0]), "replicas diverged"
            assert v["gsteps"][r].item() == step
        if dtype == torch.float32:
            # the kernel fuses w - step*s into one FMA, so allow the last ulp against torch
            assert torch.allclose(v["work"][0], expected, rtol=1e-6, atol=1e-7)
        else:
            got = torch.cat(v["masters"])
            assert torch.allclose(got, expected, atol=1e-6)
            assert torch.equal(v["work"][0], got.bfloat16())


@pytest.mark.parametrize("world", [2, 4, 8])
def test_push_sync_ll_matches_flag_protocol(native_lib, world):
    """The low-latency packet push gives bit-identical parameters to the flag-based push."""
    from b200ps import ops
    res = []
    for ll in (False, True):
        v = _virtual_ranks(world, 79510, torch.float32)
        llg = [torch.zeros(world * v["shard"], 2, device=DEV, dtype=torch.int32) for _ in range(world)]
        llp = [torch.zeros(v["total"], 2, device=DEV, dtype=torch.int32) for _ in range(world)]
        T = lambda ts: ops.device_ptr_table([t.data_ptr() for t in ts], DEV)
        t_llg, t_llp = T(llg), T(llp)
        for step in range(3):
            for g in v["grads"]:
                g.mul_(0.5).add_(0.01 * (step + 1))
            if ll:
                ops.push_sync_ll(v["t_pads"], v["t_masters"], v["t_grads"], v["bounds"], v["t_gsteps"], t_llg, t_llp,
                                 world=world, rank_base=0, local_ranks=world, max_shard_elems=v["shard"],
                                 total_elems=v["total"], lr=0.1)
            else:
                ops.push_sync_sgd(v["t_pads"], v["t_masters"], v["t_outs"], v["t_grads"], v["bounds"], v["t_gsteps"],
                                  world=world, rank_base=0, local_ranks=world, max_shard_elems=v["shard"],
                                  dtype=torch.float32, replicas_to_aggregate=world, lr=0.1)
        torch.cuda.synchronize()
        for r in range(1, world):
            assert torch.equal(v["work"][r], v["work"][0])
        assert all(gs.item() == 3 for gs in v["gsteps"])
        res.append(v["work"][0].clone())
    assert torch.equal(res[0], res[1])


def test_push_sync_is_bitwise_reproducible(native_lib):
    from b200ps import ops
    outs = []
    for _ in range(2):
        v = _virtual_ranks(4, 79510, torch.float32)
        for _step in range(3):
            ops.push_sync_sgd(v["t_pads"], v["t_masters"], v["t_outs"], v["t_grads"], v["bounds"],
                              v["t_gsteps"], world=4, rank_base=0, local_ranks=4,
                              max_shard_elems=v["shard"], dtype=torch.float32, replicas_to_aggregate=4,
                              lr=0.05)
        torch.cuda.synchronize()
        outs.append(v["work"][2].clone())
    assert torch.equal(outs[0], outs[1])


def test_push_async_and_pull_virtual_ranks(native_lib):
    from b200ps import ops
    world = 4
    v = _virtual_ranks(world, 79510, torch.float32)
    lr = 0.01
    for r in range(world):          # each worker pushes once, in any order: updates must all land
        ops.push_async(v["t_masters"], v["t_grads"], v["bounds"], v["t_gsteps"], world=world, rank_base=r,
                       total_elems=v["total"], lr=lr)
    torch.cuda.synchronize()
    want = v["init"].double()
    for r in range(world):
        want = want - lr * v["grads"][r].double()
    owners = torch.cat([v["work"][r][v["shard"] * r: v["shard"] * (r + 1)] for r in range(world)])
    assert (owners.double() - want).abs().max().item() < 1e-5        # update conservation
    assert v["gsteps"][0].item() == world                             # global_step counts every push
    ops.pull(v["t_masters"], v["t_outs"], v["bounds"], world=world, rank_base=0, local_ranks=world,
             max_shard_elems=v["shard"], dtype=torch.float32)
    torch.cuda.synchronize()
    for r in range(world):
        assert torch.equal(v["work"][r], owners)


def test_barrier_virtual_ranks(native_lib):
    from b200ps import ops
    world = 8
    pads = [torch.zeros(ops.signal_pad_words(), device=DEV, dtype=torch.int32) for _ in range(world)]
    t = ops.device_ptr_table([p.data_ptr() for p in pads], DEV)
    for _ in range(3):
        ops.barrier(t, world=world, rank_base=0, local_ranks=world)
    torch.cuda.synchronize()
    assert all(p[97].item() == 3 for p in pads)


def test_sparse_kernels_refuse_out_of_range_ids(native_lib):
    """Ids outside [0, features) never touch memory: they contribute nothing, are counted on the device and
    surface as a ValueError at the next check (embedding_lookup_sparse's InvalidArgument, lr2.py:390)."""
    from b200ps import ops
    F, B = 1000, 4                      # 2 shards x 512 rows: ids 1000..1023 fall into shard padding
    rows = 512
    shards = [torch.ones(rows, device=DEV) for _ in range(2)]
    table = ops.device_ptr_table([s.data_ptr() for s in shards], DEV)
    bias = torch.zeros(1, device=DEV)
    rp = torch.tensor([0, 2, 4, 6, 8], device=DEV, dtype=torch.int32)
    fid = torch.tensor([1, 2, 999, 1000, 5, -7, 6, 1 << 40], device=DEV, dtype=torch.int64)
    fv = torch.ones(8, device=DEV)
    y = torch.zeros(B, device=DEV)
    pred = torch.empty(B, device=DEV); loss = torch.zeros(1, device=DEV)
    ops.sparse_bad_ids(reset=True)
    ops.sparse_lr(rp, fid, fv, y, table, bias, rows_per_shard=rows, features=F, pred=pred, loss_sum=loss,
                  forward=True, update=True, lr=1.0)
    torch.cuda.synchronize()
    z = torch.tensor([2.0, 1.0, 1.0, 1.0], device=DEV)      # the refused non-zeros contribute nothing
    assert torch.allclose(pred, torch.sigmoid(z), atol=1e-6)
    assert torch.cat(shards)[1000:].eq(1.0).all()            # padding rows untouched
    assert ops.sparse_bad_ids(reset=False) == 3
    with pytest.raises(ValueError):
        ops.check_sparse_ids()
    ops.check_sparse_ids()                                   # counter was cleared


@pytest.mark.parametrize("B,D,H,Cn", [(100, 784, 100, 10), (37, 48, 20, 3), (128, 64, 128, 16), (8, 32, 4, 2),
                                      (100, 1024, 128, 16)])
def test_mlp_cluster_steps_match_float64_reference(native_lib, B, D, H, Cn):
    """K14 (resident 16-CTA cluster, tensor-core products with the 3xTF32 split, DSMEM exchanges): three steps
    in ONE launch over a resident dataset == three float64 SGD steps; one launch of 3 == three launches of 1
    bit for bit; plain-tf32 mode stays within tf32 rounding."""
    from b200ps import ops
    if not ops.mlp_cluster_supported(B, D, H, Cn):
        pytest.skip("16-CTA clusters with ~200 KB shared memory are not schedulable on this device")
    g = torch.Generator(device=DEV).manual_seed(4)
    rows = 4 * B
    x = torch.rand(rows, D, device=DEV, generator=g)
    labels = torch.randint(0, Cn, (rows,), device=DEV, generator=g, dtype=torch.int32)
    init = [torch.randn(D, H, device=DEV, generator=g) * 0.2, torch.randn(H, device=DEV, generator=g) * 0.1,
            torch.randn(H, Cn, device=DEV, generator=g) * 0.5, torch.randn(Cn, device=DEV, generator=g) * 0.1]
    lr, steps, start = 0.1, 3, 2 * B

    # float64 oracle: rows [2B,3B), [3B,4B), then wrap to [0,B)
    W1, B1, W2, B2 = (t.double().clone() for t in init)
    ref_loss, ref_corr = [], []
    for s in range(steps):
        r0 = (start + s * B) % rows
        xb, yb = x[r0:r0 + B].double(), labels[r0:r0 + B].long()
        ps = [t.clone().requires_grad_(True) for t in (W1, B1, W2, B2)]
        logits = torch.sigmoid(xb @ ps[0] + ps[1]) @ ps[2] + ps[3]
        loss = torch.nn.functional.cross_entropy(logits, yb)
        loss.backward()
        ref_loss.append(loss.item()); ref_corr.append(int((logits.argmax(1) == yb).sum()))
        W1, B1, W2, B2 = (t - lr * q.grad for t, q in zip((W1, B1, W2, B2), ps))

    def run(nlaunch, plain=False):
        w = [t.clone() for t in init]
        cursor = torch.tensor([start], device=DEV, dtype=torch.int32)
        gstep = torch.zeros(1, device=DEV, dtype=torch.int64)
        loss = torch.zeros(steps, 2, device=DEV); corr = torch.zeros(steps, 2, device=DEV, dtype=torch.int32)
        per = steps // nlaunch
        for l in range(nlaunch):
            ops.mlp_cluster_steps(x, labels, *w, loss=loss[l * per:, 0], correct=corr[l * per:, 1], stat_stride=2,
                                  steps=per, lr=lr, batch=B, cursor=cursor, global_step=gstep, cursor_rw=cursor,
                                  cursor_step=B, cursor_wrap=rows, plain_tf32=plain)
        torch.cuda.synchronize()
        return w, loss[:, 0].tolist(), corr[:, 1].tolist(), cursor.item(), gstep.item()

    w, losses, corrs, cur, gs = run(1)
    assert cur == (start + steps * B) % rows and gs == steps
    for got, want in zip(losses, ref_loss):
        assert abs(got - want) < 2e-5 * max(1.0, abs(want)), (losses, ref_loss)
    assert corrs == ref_corr
    for t, want in zip(w, (W1, B1, W2, B2)):
        assert torch.allclose(t.double(), want, atol=5e-6, rtol=2e-5), (t.double() - want).abs().max()
    w3, losses3, corrs3, _, _ = run(3)                       # three launches of one step: same bytes
    for a, b in zip(w, w3):
        assert torch.equal(a, b)
    assert losses == losses3 and corrs == corrs3
    wp, lossp, _, _, _ = run(1, plain=True)                  # one tf32 product per tile
    for t, want in zip(wp, (W1, B1, W2, B2)):
        assert torch.allclose(t.double(), want, atol=5e-3, rtol=5e-3)
    assert abs(lossp[0] - ref_loss[0]) < 5e-3 * max(1.0, abs(ref_loss[0]))


def test_workspace_refuses_overlapping_streams(native_lib):
    """The library workspace is one block per device: an eager launch that needs it while a launch on ANOTHER stream
    may still be using it is refused (instead of two kernels sharing scratch); the same stream, or the other stream
    once the first has finished, is fine."""
    from b200ps import ops
    from b200ps.ops.native import NativeError
    B, H, C = 64, 32, 10
    g = torch.Generator(device=DEV).manual_seed(2)
    z2 = torch.randn(B, H, device=DEV, generator=g); b1 = torch.zeros(H, device=DEV)
    w2 = torch.randn(H, C, device=DEV, generator=g); b2 = torch.zeros(C, device=DEV)
    labels = torch.randint(0, C, (B,), device=DEV, generator=g, dtype=torch.int32)
    loss = torch.zeros(1, device=DEV); correct = torch.zeros(1, device=DEV, dtype=torch.int32)
    kw = dict(loss_sum=loss, correct=correct, train=False, loss_scale=1.0 / B)
    sa, sb = torch.cuda.Stream(), torch.cuda.Stream()
    torch.cuda.synchronize()
    with torch.cuda.stream(sa):
        torch.cuda._sleep(400_000_000)                      # ~0.2 s of GPU time queued ahead on stream A
        ops.mlp_tail(z2, b1, w2, b2, labels, **kw)
        ops.mlp_tail(z2, b1, w2, b2, labels, **kw)          # same stream: ordered, allowed
    with torch.cuda.stream(sb):
        with pytest.raises(NativeError, match="another stream"):
            ops.mlp_tail(z2, b1, w2, b2, labels, **kw)
    torch.cuda.synchronize()
    with torch.cuda.stream(sb):
        ops.mlp_tail(z2, b1, w2, b2, labels, **kw)          # stream A has finished: fine
    torch.cuda.synchronize()


@pytest.mark.parametrize("world,dim,wide,sync", [(4, 128, True, True), (2, 256, False, False), (8, 128, True, False)])
def test_wide_deep_owner_side_step_matches_float64_reference(native_lib, world, dim, wide, sync):
    """K4o/K5o on virtual ranks (all ranks of each phase on one GPU): request / serve / loss / fold rounds leave the
    tables where the float64 oracle puts them — sync: mean of the replicas' gradients; async rounds: their sum."""
    from b200ps import ops
    from b200ps.models.wide_deep import SparseWideDeep, wide_deep_reference
    from b200ps.parallel.context import init_from_env
    from b200ps.ps.store import ParameterStore
    ctx = init_from_env(virtual_world=world)
    F, B = 1 << 13, 120
    store = ParameterStore(ctx.pool)
    model = SparseWideDeep(F, dim=dim, wide=wide, stddev=0.1, owner=True)
    model.register(store)
    store.build(); store.initialize()
    init = {k: v.clone() for k, v in store.state_dict().items()}
    batches = {r: _sparse_batch(B=B, nnz=30, spread=8, F=F, seed=40 + r) for r in range(world)}
    lr = 0.5
    losses = model._step_owner(store, batches, lr, sync=sync)
    torch.cuda.synchronize()
    got = store.state_dict()
    W0 = init[model.W].reshape(-1) if wide else None
    E0, w20, b0 = init[model.E], init[model.W2].reshape(-1), init[model.Bn].reshape(())
    acc = {"E": torch.zeros_like(E0, dtype=torch.float64), "w2": torch.zeros(dim, dtype=torch.float64), "b": 0.0,
           "W": torch.zeros(F, dtype=torch.float64)}
    for r in range(world):
        rp, fid, fv, y = (t.cpu() for t in batches[r])
        loss_ref, _, gr = wide_deep_reference(W0, E0, w20, b0, rp, fid, fv, y)
        assert abs(losses[r].item() - loss_ref.item()) < 2e-5
        for k in gr:
            acc[k] = acc[k] + gr[k]
    scale = lr / world if sync else lr
    assert (got[model.E].double() - (E0.double() - scale * acc["E"])).abs().max().item() < 2e-5
    assert (got[model.W2].reshape(-1).double() - (w20.double() - scale * acc["w2"])).abs().max().item() < 5e-5
    if wide:
        assert (got[model.W].reshape(-1).double() - (W0.double() - scale * acc["W"])).abs().max().item() < 2e-5
    assert abs(got[model.Bn].item() - (b0.item() - scale * float(acc["b"]))) < 2e-5
    assert store.sparse_shard(model.P, 0).abs().sum().item() == 0.0          # pending table folded and cleared
    assert store.global_step == (1 if sync else world)
    # a second round from the updated tables still agrees with the oracle's loss at those tables
    losses2 = model._step_owner(store, batches, lr, sync=sync)
    torch.cuda.synchronize()
    rp, fid, fv, y = (t.cpu() for t in batches[0])
    l2, _, _ = wide_deep_reference(got[model.W].reshape(-1) if wide else None, got[model.E], got[model.W2].reshape(-1),
                                   got[model.Bn].reshape(()), rp, fid, fv, y, train=False)
    assert abs(losses2[0].item() - l2.item()) < 2e-5
