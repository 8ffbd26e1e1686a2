"""Metric and model math restated on the CPU (P11, P15, P16, P27) vs independent oracles."""
import math

import numpy as np
import pytest
import torch

from b200ps.metrics.auc import StreamingAUC, auc_from_counts, auc_from_counts_reference, make_thresholds
from b200ps.models.linreg import LinearRegression, make_data
from b200ps.models.mlp import mlp_reference_step
from b200ps.models.sparse_lr import sparse_lr_reference
from b200ps.optim.adam import Adam


def test_auc_thresholds_and_counts_match_oracle():
    thr = make_thresholds(200)
    assert len(thr) == 200 and thr[0] < 0 < thr[1] and thr[-1] > 1 and thr[100] == np.float32(100 / 199)
    rng = np.random.RandomState(0)
    pred = rng.rand(3000).astype(np.float32)
    pred[:6] = [0.0, 1.0, 1 / 199, 50 / 199, 198 / 199, 0.5]               # exactly on thresholds
    y = (rng.rand(3000) < pred).astype(np.float32)
    m = StreamingAUC(200)
    m.update(torch.from_numpy(pred[:1000]), torch.from_numpy(y[:1000]))
    m.update(torch.from_numpy(pred[1000:]), torch.from_numpy(y[1000:]))      # streaming: accumulates
    for got, want in zip(m.counts(), auc_from_counts_reference(pred, y, 200)):
        assert np.array_equal(got, want)
    # float64 trapezoid vs exact pairwise AUC: the 200-bucket approximation is close
    pos, neg = pred[y > 0.5], pred[y <= 0.5]
    exact = ((pos[:, None] > neg[None, :]).mean() + 0.5 * (pos[:, None] == neg[None, :]).mean())
    assert abs(m.result() - exact) < 5e-3
    assert auc_from_counts(*m.counts()) == m.result()
    m.reset(); assert m.counts()[0].sum() == 0


def test_mlp_reference_matches_autograd():
    g = torch.Generator().manual_seed(0)
    names = ("weights/Variable", "weights/Variable_1", "biases/Variable", "biases/Variable_1")
    p = {names[0]: torch.randn(784, 100, generator=g) * 0.1, names[1]: torch.randn(100, 10, generator=g) * 0.1,
         names[2]: torch.randn(100, generator=g) * 0.1, names[3]: torch.randn(10, generator=g) * 0.1}
    x = torch.rand(100, 784, generator=g); y = torch.randint(0, 10, (100,), generator=g)
    pr = {k: v.clone().requires_grad_(True) for k, v in p.items()}
    z = torch.sigmoid(x @ pr[names[0]] + pr[names[2]]) @ pr[names[1]] + pr[names[3]]
    # the reference's loss: mean(-sum(y_ * log(softmax(z)))) (example.py:98,103)
    onehot = torch.nn.functional.one_hot(y, 10).float()
    loss = torch.mean(-torch.sum(onehot * torch.log(torch.softmax(z, 1)), dim=1))
    loss.backward()
    l2, grads, correct = mlp_reference_step(p, x, y)
    assert abs(l2.item() - loss.item()) < 1e-5 and correct == int((z.argmax(1) == y).sum())
    for k in names:
        assert torch.allclose(grads[k], pr[k].grad, atol=1e-6)


def test_sparse_lr_reference_matches_autograd():
    g = torch.Generator().manual_seed(1)
    F, B = 50, 8
    lens = torch.randint(1, 6, (B,), generator=g)
    rp = torch.zeros(B + 1, dtype=torch.int32); rp[1:] = torch.cumsum(lens, 0)
    n = int(rp[-1])
    fid = torch.randint(0, F, (n,), generator=g); fv = torch.rand(n, generator=g)
    y = torch.randint(0, 2, (B,), generator=g).float()
    W = torch.randn(F, 1, generator=g, dtype=torch.float64, requires_grad=True)
    b = torch.zeros(1, dtype=torch.float64, requires_grad=True)
    rows = torch.repeat_interleave(torch.arange(B), lens)
    z = torch.zeros(B, dtype=torch.float64).index_add(0, rows, W[fid, 0] * fv.double()) + b
    loss = torch.nn.functional.binary_cross_entropy_with_logits(z, y.double())     # tf sigmoid_cross_entropy mean
    loss.backward()
    l2, pred, (ids, vals), db = sparse_lr_reference(W.detach(), b.detach(), rp, fid, fv, y)
    assert abs(float(l2) - float(loss)) < 1e-12 and torch.allclose(pred, torch.sigmoid(z.detach()))
    dW = torch.zeros(F, dtype=torch.float64).index_add_(0, ids, vals)
    assert torch.allclose(dW, W.grad[:, 0]) and abs(float(db) - float(b.grad)) < 1e-12


def test_adam_matches_tf_formula_and_slot_names():
    opt = Adam(0.01)
    w = torch.tensor([[0.5]]); b = torch.zeros(1)
    wr, mr, vr = 0.5, 0.0, 0.0
    for t in range(1, 6):
        g = torch.tensor([[0.3 * t]])
        opt.apply({"test/weights": w, "test/bias": b}, {"test/weights": g, "test/bias": torch.ones(1)})
        gd = 0.3 * t
        mr = 0.9 * mr + 0.1 * gd; vr = 0.999 * vr + 0.001 * gd * gd
        wr -= 0.01 * math.sqrt(1 - 0.999 ** t) / (1 - 0.9 ** t) * mr / (math.sqrt(vr) + 1e-8)
    assert abs(w.item() - wr) < 1e-6
    slots = opt.slot_variables("test/")
    assert set(slots) == {"test/test/weights/Adam", "test/test/weights/Adam_1", "test/test/bias/Adam",
                          "test/test/bias/Adam_1", "test/beta1_power", "test/beta2_power"}
    assert slots["test/beta1_power"].item() == pytest.approx(0.9 ** 6)


def test_linear_regression_demo_converges_cpu():
    x, y = make_data(1000)
    xd, yd = torch.from_numpy(x), torch.from_numpy(y)
    m = LinearRegression("cpu", seed=0)
    rng = np.random.RandomState(0)
    first = None
    for i in range(500):
        idx = torch.from_numpy(rng.choice(1000, 100))
        loss = m.train_step(xd[idx], yd[idx])
        first = float(loss) if first is None else first
    assert float(loss) < first and 0.8 < m.w.item() < 1.4       # y ~ x + bounded wiggle => slope near 1
    v = m.variables()
    assert v["test/weights"].shape == (1, 1) and v["test/bias"].shape == (1,)


def test_wide_deep_reference_matches_autograd():
    """The float64 oracle of the wide & deep model (lr2_debug.py:423-428) against torch autograd."""
    import torch
    from b200ps.models.wide_deep import wide_deep_reference
    g = torch.Generator().manual_seed(3)
    F, D, B = 50, 8, 6
    lens = torch.randint(1, 5, (B,), generator=g)
    row_ptr = torch.cat([torch.zeros(1, dtype=torch.long), lens.cumsum(0)]).int()
    n = int(row_ptr[-1])
    fid = torch.randint(0, F, (n,), generator=g)
    fval = torch.rand(n, generator=g)
    y = (torch.rand(B, generator=g) > 0.5).float()
    W = torch.randn(F, 1, generator=g, dtype=torch.float64, requires_grad=True)
    E = torch.randn(F, D, generator=g, dtype=torch.float64, requires_grad=True)
    w2 = torch.randn(D, 1, generator=g, dtype=torch.float64, requires_grad=True)
    b = torch.zeros(1, dtype=torch.float64, requires_grad=True)
    row_of = torch.repeat_interleave(torch.arange(B), lens)
    z = torch.zeros(B, dtype=torch.float64).index_add(0, row_of, W[fid, 0] * fval) \
        + (torch.zeros(B, D, dtype=torch.float64).index_add(0, row_of, E[fid] * fval[:, None]) @ w2)[:, 0] + b
    loss = torch.nn.functional.binary_cross_entropy_with_logits(z, y.double())
    loss.backward()
    l2, pred, gr = wide_deep_reference(W.detach(), E.detach(), w2.detach(), b.detach(), row_ptr, fid, fval, y)
    assert abs(l2.item() - loss.item()) < 1e-12 and torch.allclose(pred, torch.sigmoid(z.detach()))
    assert torch.allclose(gr["W"], W.grad[:, 0]) and torch.allclose(gr["E"], E.grad)
    assert torch.allclose(gr["w2"], w2.grad[:, 0]) and torch.allclose(gr["b"], b.grad[0])


def test_accuracy_metric_one_hot_and_streaming():
    import torch
    from b200ps.metrics import StreamingAccuracy, accuracy
    logits = torch.tensor([[0.1, 0.9, 0.0], [0.5, 0.5, 0.1], [0.0, 0.2, 0.7], [1.0, 0.0, 0.0]])
    labels = torch.tensor([1, 0, 1, 0])
    assert accuracy(logits, labels) == 0.75                                  # tie -> lowest index (row 1 correct)
    assert accuracy(logits, torch.nn.functional.one_hot(labels, 3)) == 0.75  # the reference feeds one-hot y_
    m = StreamingAccuracy()
    m.update_from_logits(logits[:2], labels[:2]); m.update(torch.tensor([1], dtype=torch.int32), 2)
    assert m.result() == 0.75
    m.reset(); m.update(3, 4)
    assert m.result() == 0.75
