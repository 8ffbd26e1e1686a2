"""Torch-tensor front ends of the native sm_100a kernels (libb200ps.so).

Each function validates shapes/dtypes/devices, then hands raw device pointers and the
current CUDA stream to the C ABI (include/b200ps.h).  There is no PyTorch fallback: on a
CUDA tensor the native kernel runs or an exception is raised.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import torch

from . import native
from .native import (ACT_NONE, ACT_SIGMOID, DT_BF16, DT_F32, EPI_ATOMIC_ADD, EPI_SGD, EPI_STORE,
                     GemmArgs, check, ptr, stream_ptr)

__all__ = [
    "gemm", "mlp_tail", "mlp_fused_step", "mlp_cluster_steps", "mlp_cluster_supported", "softmax_ce", "colsum", "sgd_apply", "adam_apply", "adam_advance",
    "linreg_fwd_bwd", "synth_mnist", "synth_dense_bf16", "synth_sparse", "init_normal",
    "cast_f32_bf16", "sparse_lr", "sparse_lr_unique", "sparse_dedup_table_size", "sparse_dedup_list_size", "sparse_bad_ids", "check_sparse_ids", "sparse_wd", "sparse_wd_apply", "wd_owner_request", "wd_owner_serve", "wd_owner_loss", "wd_owner_fold", "auc_hist", "push_sync_sgd", "push_sync_ll", "store_nvls", "push_async", "pull", "barrier",
    "device_ptr_table", "signal_pad_words", "copy_words",
]


def _dt(t: torch.Tensor) -> int:
    if t.dtype == torch.bfloat16:
        return DT_BF16
    if t.dtype == torch.float32:
        return DT_F32
    raise TypeError(f"unsupported dtype {t.dtype} (bf16 or fp32 only)")


def _cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise ValueError("b200ps kernels need CUDA tensors")


def _row_stride(t: torch.Tensor) -> int:
    if t.dim() != 2 or t.stride(1) != 1:
        raise ValueError("expected a 2-D tensor with unit inner stride")
    return t.stride(0)


def gemm(a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None, *,
         a_mn_major: bool = False, b_mn_major: bool = False, alpha: float = 1.0,
         bias: Optional[torch.Tensor] = None, act: int = ACT_NONE,
         dsigmoid_of: Optional[torch.Tensor] = None, epilogue: int = EPI_STORE, split_k: int = 1,
         sgd_master: Optional[torch.Tensor] = None, sgd_lowp: Optional[torch.Tensor] = None,
         sgd_lr: float = 0.0, out_dtype: Optional[torch.dtype] = None, max_ctas: int = 0,
         a_row_offset: Optional[torch.Tensor] = None, a_rows: Optional[int] = None,
         c_split_stride: int = 0, gather: Optional[dict] = None,
         colsum_out: Optional[torch.Tensor] = None, stream=None) -> Optional[torch.Tensor]:
    """tcgen05/TMA GEMM  ``out[M,N] = epilogue(alpha * op(a) @ op(b))``.

    ``a`` is ``[M,K]`` (K-major) or, with ``a_mn_major``, stored ``[K,M]``;
    ``b`` is ``[N,K]`` (K-major) or, with ``b_mn_major``, stored ``[K,N]``.
    fp32 operands run on the tensor cores as tf32.  Replaces tf.matmul and its gradients
    (example.py:95,97,118).

    Resident datasets: with ``a_row_offset`` (int32 device scalar) and ``a_rows`` (rows this call
    consumes), ``a`` is the WHOLE stored array and the kernel reads rows
    ``[*a_row_offset, *a_row_offset + a_rows)`` of it — the offset is dereferenced on the device,
    so a CUDA graph can walk a dataset without per-step copies.
    """
    _cuda(a, b, out, bias, dsigmoid_of, sgd_master, sgd_lowp)
    if a.dtype != b.dtype:
        raise TypeError("gemm: A and B dtypes differ")
    M, K = (a.shape[1], a.shape[0]) if a_mn_major else (a.shape[0], a.shape[1])
    a_rows_total = 0
    if a_row_offset is not None:
        if a_rows is None or a_row_offset.dtype != torch.int32:
            raise ValueError("gemm: a_row_offset needs a_rows and an int32 device scalar")
        a_rows_total = a.shape[0]
        if a_mn_major:
            K = a_rows
        else:
            M = a_rows
    N, Kb = (b.shape[1], b.shape[0]) if b_mn_major else (b.shape[0], b.shape[1])
    if K != Kb:
        raise ValueError(f"gemm: reduction dims differ ({K} vs {Kb})")
    args = GemmArgs()
    args.M, args.N, args.K = M, N, K
    args.ab_dtype = _dt(a)
    args.a_mn_major, args.b_mn_major = int(a_mn_major), int(b_mn_major)
    args.A, args.lda = ptr(a), _row_stride(a)
    args.B, args.ldb = ptr(b), _row_stride(b)
    if epilogue == EPI_SGD:
        if sgd_master is None or tuple(sgd_master.shape) != (M, N) or sgd_master.dtype != torch.float32:
            raise ValueError("gemm: EPI_SGD needs an fp32 sgd_master of shape [M,N]")
        args.C, args.ldc, args.c_dtype = 0, _row_stride(sgd_master), DT_F32
        if sgd_lowp is not None and (_row_stride(sgd_lowp) != args.ldc or sgd_lowp.dtype != torch.bfloat16):
            raise ValueError("gemm: sgd_lowp must be bf16 with the master's row stride")
    else:
        if out is None:
            out = torch.empty((M, N), device=a.device,
                              dtype=out_dtype or (torch.float32 if epilogue == EPI_ATOMIC_ADD else a.dtype))
        if tuple(out.shape) != (M, N):
            raise ValueError(f"gemm: out has shape {tuple(out.shape)}, want {(M, N)}")
        args.C, args.ldc, args.c_dtype = ptr(out), _row_stride(out), _dt(out)
    args.epilogue = epilogue
    args.alpha = alpha
    if bias is not None and (bias.dtype != torch.float32 or bias.numel() != N):
        raise ValueError("gemm: bias must be fp32 [N]")
    args.bias = ptr(bias)
    args.act = act
    if dsigmoid_of is not None:
        if out is None or dsigmoid_of.dtype != out.dtype or tuple(dsigmoid_of.shape) != (M, N):
            raise ValueError("gemm: dsigmoid_of must match out")
        args.aux, args.ldaux = ptr(dsigmoid_of), _row_stride(dsigmoid_of)
    else:
        args.aux, args.ldaux = 0, 0
    args.split_k = split_k
    args.sgd_master, args.sgd_lowp, args.sgd_lr = ptr(sgd_master), ptr(sgd_lowp), sgd_lr
    args.max_ctas = max_ctas
    args.a_row_offset, args.a_rows_total = ptr(a_row_offset), a_rows_total
    args.c_split_stride = c_split_stride
    args.colsum_out = ptr(colsum_out)          # fused bias gradient (column sums of the stored tile)
    if gather is not None:
        # K2: pull fused into the GEMM — gather = dict(src_table, flags, world, rank[, ctas])
        args.gather_src, args.gather_flags = ptr(gather["src_table"]), ptr(gather["flags"])
        args.gather_world, args.gather_rank = gather["world"], gather["rank"]
        args.gather_ctas = gather.get("ctas", 0)
    check(native.load().b200ps_gemm(C.byref(args), stream_ptr(stream)), "gemm")
    return out


def mlp_tail(z2, b1, w2, b2, labels, *, loss_sum, correct=None, dz2=None, dw2=None, db2=None,
             db1=None, a2_out=None, probs_out=None, train=True, grad_scale=None, loss_scale=None,
             overwrite_stats=False, clear_z2=False, label_offset=None, z2_splits=1, z2_split_stride=0,
             stream=None):
    """Fused narrow MLP head (K6): see include/b200ps.h b200ps_mlp_tail."""
    _cuda(z2, b1, w2, b2, labels)
    B, H = z2.shape
    Cn = w2.shape[1]
    if labels.dtype != torch.int32:
        raise TypeError("mlp_tail: labels must be int32 class ids")
    if label_offset is not None and label_offset.dtype != torch.int32:
        raise TypeError("mlp_tail: label_offset must be an int32 device scalar")
    gs = (1.0 / B) if grad_scale is None else grad_scale
    ls = (1.0 / B) if loss_scale is None else loss_scale
    check(native.load().b200ps_mlp_tail(ptr(z2), ptr(b1), ptr(w2), ptr(b2), ptr(labels), ptr(a2_out),
                                        ptr(dz2), ptr(dw2), ptr(db2), ptr(db1), ptr(loss_sum),
                                        ptr(correct), ptr(probs_out), B, H, Cn, int(train), gs, ls,
                                        (native.TAIL_OVERWRITE_STATS if overwrite_stats else 0) |
                                        (native.TAIL_CLEAR_Z2 if clear_z2 else 0), ptr(label_offset), z2_splits,
                                        z2_split_stride,
                                        stream_ptr(stream)), "mlp_tail")


def mlp_fused_step(x, labels, w1, b1, w2, b2, *, loss, correct=None, grads=None, lr=0.0, batch=None, cursor=None,
                   global_step=None, cursor_rw=None, cursor_step=0, cursor_wrap=1, grad_scale=None, loss_scale=None,
                   push=None, hogwild=False, stream=None):
    """Whole narrow-MLP training step in one launch (K13): see include/b200ps.h b200ps_mlp_fused_step.
    ``grads=None``: apply SGD in place (single replica); ``grads=(gw1, gb1, gw2, gb2)``: write gradients;
    ``push=native.FusedPush(...)``: synchronous-replica push inside the launch (one rank per GPU);
    with ``hogwild=True`` the same tables serve the asynchronous in-kernel step (no local parameter copy)."""
    _cuda(x, labels, w1, b1, w2, b2)
    B = x.shape[0] if batch is None else batch
    D, H = w1.shape
    Cn = w2.shape[1]
    if labels.dtype != torch.int32:
        raise TypeError("mlp_fused_step: labels must be int32 class ids")
    gs = (1.0 / B) if grad_scale is None else grad_scale
    ls = (1.0 / B) if loss_scale is None else loss_scale
    g = grads or (None, None, None, None)
    mode = (3 if hogwild else 2) if push is not None else int(grads is None)
    import ctypes
    check(native.load().b200ps_mlp_fused_step(
        ptr(x), ptr(labels), ptr(cursor), ptr(w1), ptr(b1), ptr(w2), ptr(b2), ptr(g[0]), ptr(g[1]), ptr(g[2]),
        ptr(g[3]), ptr(loss), ptr(correct), ptr(global_step), ptr(cursor_rw), cursor_step, cursor_wrap, B, D, H, Cn,
        mode, lr, gs, ls, ctypes.addressof(push) if push is not None else None, stream_ptr(stream)), "mlp_fused_step")


def mlp_cluster_supported(batch: int, in_dim: int, hidden: int, classes: int) -> bool:
    """Does the resident-cluster step kernel (K14, csrc/kernels/mlp_cluster.cu) cover this shape on this device?"""
    ok = C.c_int32(0)
    check(native.load().b200ps_mlp_cluster_supported(batch, in_dim, hidden, classes, C.byref(ok)), "mlp_cluster_supported")
    return bool(ok.value)


def mlp_cluster_steps(x, labels, w1, b1, w2, b2, *, loss, correct=None, stat_stride=0, steps=1, lr, batch=None,
                      cursor=None, global_step=None, cursor_rw=None, cursor_step=None, cursor_wrap=None,
                      grad_scale=None, loss_scale=None, plain_tf32=False, stream=None):
    """``steps`` consecutive single-replica training steps of the narrow MLP in ONE launch of one 16-CTA cluster
    (K14): step s reads rows ``[(cursor + s*cursor_step) % cursor_wrap, +batch)`` of ``x`` / ``labels`` and writes
    its loss / correct count to ``loss[s*stat_stride]`` / ``correct[s*stat_stride]``; parameters are updated in
    place (SGD), ``global_step += steps``, ``cursor_rw`` receives the advanced cursor."""
    _cuda(x, labels, w1, b1, w2, b2)
    B = x.shape[0] if batch is None else batch
    D, H = w1.shape
    Cn = w2.shape[1]
    if labels.dtype != torch.int32:
        raise TypeError("mlp_cluster_steps: labels must be int32 class ids")
    gs = (1.0 / B) if grad_scale is None else grad_scale
    ls = (1.0 / B) if loss_scale is None else loss_scale
    cstep = B if cursor_step is None else cursor_step
    cwrap = max(x.shape[0], steps * cstep) if cursor_wrap is None else cursor_wrap
    check(native.load().b200ps_mlp_cluster_steps(
        ptr(x), ptr(labels), ptr(cursor), ptr(w1), ptr(b1), ptr(w2), ptr(b2), ptr(loss), ptr(correct), stat_stride,
        ptr(global_step), ptr(cursor_rw), cstep, cwrap, B, D, H, Cn, steps, lr, gs, ls, int(plain_tf32),
        stream_ptr(stream)), "mlp_cluster_steps")


def softmax_ce(logits, labels, *, dlogits=None, loss_sum=None, correct=None, grad_scale=None,
               loss_scale=None, stream=None):
    _cuda(logits, labels)
    B, Cn = logits.shape
    gs = (1.0 / B) if grad_scale is None else grad_scale
    ls = (1.0 / B) if loss_scale is None else loss_scale
    check(native.load().b200ps_softmax_ce(ptr(logits), ptr(labels), ptr(dlogits), ptr(loss_sum),
                                          ptr(correct), B, Cn, _row_stride(logits), _dt(logits), gs, ls,
                                          stream_ptr(stream)), "softmax_ce")


def colsum(x, out, scale=1.0, stream=None):
    _cuda(x, out)
    M, N = x.shape
    check(native.load().b200ps_colsum(ptr(x), ptr(out), M, N, _row_stride(x), _dt(x), scale,
                                      stream_ptr(stream)), "colsum")


def sgd_apply(params, grads, lr, *, lowp=None, grad_scale=1.0, stream=None):
    _cuda(params, grads, lowp)
    check(native.load().b200ps_sgd_apply(ptr(params), ptr(grads), _dt(grads), ptr(lowp), params.numel(),
                                         lr, grad_scale, stream_ptr(stream)), "sgd_apply")


def adam_apply(params, grads, m, v, beta_pow, lr, beta1=0.9, beta2=0.999, eps=1e-8, stream=None):
    _cuda(params, grads, m, v, beta_pow)
    check(native.load().b200ps_adam_apply(ptr(params), ptr(grads), ptr(m), ptr(v), params.numel(), lr,
                                          beta1, beta2, eps, ptr(beta_pow), stream_ptr(stream)),
          "adam_apply")


def adam_advance(beta_pow, beta1=0.9, beta2=0.999, stream=None):
    check(native.load().b200ps_adam_advance(ptr(beta_pow), beta1, beta2, stream_ptr(stream)), "adam_advance")


def linreg_fwd_bwd(x, y, w, b, *, ypred=None, loss=None, gw=None, gb=None, inv_n=None, stream=None):
    _cuda(x, y, w, b)
    B, D = x.shape
    check(native.load().b200ps_linreg_fwd_bwd(ptr(x), ptr(y), ptr(w), ptr(b), ptr(ypred), ptr(loss),
                                              ptr(gw), ptr(gb), B, D,
                                              (1.0 / B) if inv_n is None else inv_n,
                                              stream_ptr(stream)), "linreg_fwd_bwd")


def synth_mnist(x, labels, *, onehot=None, classes=10, seed=1, offset=0, stream=None):
    _cuda(x, labels, onehot)
    B, D = x.shape
    check(native.load().b200ps_synth_mnist(ptr(x), ptr(onehot), ptr(labels), B, D, classes, seed, offset,
                                           stream_ptr(stream)), "synth_mnist")


def synth_dense_bf16(x, labels, *, classes, seed=1, offset=0, stream=None):
    _cuda(x, labels)
    B, D = x.shape
    check(native.load().b200ps_synth_dense_bf16(ptr(x), ptr(labels), B, D, classes, seed, offset,
                                                stream_ptr(stream)), "synth_dense_bf16")


def synth_sparse(row_ptr, fid, fval, labels, *, nnz_mean, spread, features, skew=1.0, seed=1, offset=0,
                 stream=None):
    _cuda(row_ptr, fid, fval, labels)
    B = labels.numel()
    if fid.numel() < B * (nnz_mean + spread):
        raise ValueError("synth_sparse: fid/fval too small for B*(nnz_mean+spread)")
    check(native.load().b200ps_synth_sparse(ptr(row_ptr), ptr(fid), ptr(fval), ptr(labels), B, nnz_mean,
                                            spread, features, skew, seed, offset, stream_ptr(stream)),
          "synth_sparse")


def init_normal(out, *, mean=0.0, stddev=1.0, seed=1, stream_id=0, elem_offset=0, stream=None):
    _cuda(out)
    if out.dtype != torch.float32 or not out.is_contiguous():
        raise TypeError("init_normal: contiguous fp32 only")
    check(native.load().b200ps_init_normal(ptr(out), out.numel(), mean, stddev, seed, stream_id,
                                           elem_offset, stream_ptr(stream)), "init_normal")


def cast_f32_bf16(src, dst, stream=None):
    _cuda(src, dst)
    check(native.load().b200ps_cast_f32_bf16(ptr(src), ptr(dst), src.numel(), stream_ptr(stream)), "cast")


def device_ptr_table(ptrs: Sequence[int], device) -> torch.Tensor:
    """int64 device tensor holding raw addresses (the pointer tables the comm kernels take)."""
    return torch.tensor([int(p) for p in ptrs], dtype=torch.int64, device=device)


def store_nvls(pads, mc_grads: int, mc_work: int, work_local: int, master: int, lo: int, hi: int, *, dtype, world: int,
               rank: int, lr: float, reduce: bool, broadcast: bool = True, global_step: Optional[int] = None, cursor=None,
               cursor_step: int = 0, cursor_wrap: int = 1, stream=None) -> None:
    """The ParameterStore's NVLS push (``reduce``) / collective pull through NVSwitch multicast: raw multicast and
    unicast ADDRESSES of buffers from ``SymmPool.alloc(..., multicast=True)`` (csrc/kernels/ps_nvls.cu)."""
    check(native.load().b200ps_store_nvls(ptr(pads), mc_grads or None, mc_work, work_local, master, lo, hi,
                                          DT_F32 if dtype == torch.float32 else DT_BF16, world, rank, lr, int(reduce),
                                          int(broadcast), global_step or None, ptr(cursor) or None, cursor_step, cursor_wrap,
                                          stream_ptr(stream)), "store_nvls")


def sparse_bad_ids(reset: bool = True) -> int:
    """Number of feature ids the sparse kernels refused on this device since the last reset (ids outside
    ``[0, features)`` never touch memory; they are skipped and counted).  Synchronises the device."""
    n = C.c_uint64()
    check(native.load().b200ps_sparse_bad_ids(C.byref(n), int(reset)), "sparse_bad_ids")
    return int(n.value)


def check_sparse_ids() -> None:
    """Raise ``ValueError`` if any step since the last check saw a feature id outside ``[0, features)`` — the
    InvalidArgument of ``embedding_lookup_sparse`` (lr2.py:390), reported at the caller's next sync point."""
    n = sparse_bad_ids(reset=True)
    if n:
        raise ValueError(f"{n} feature id(s) outside [0, features) were skipped by the sparse kernels "
                         f"(bad input line or wrong --features)")


def sparse_lr(row_ptr, fid, fval, labels, w_table, bias, *, rows_per_shard, features=None, pred=None, dz_out=None,
              dz_in=None, loss_sum=None, global_step=None, forward=True, update=True, lr=0.0,
              grad_scale=None, loss_scale=None, stream=None):
    """Sparse LR step over a row-sharded weight (K4/K5).  ``w_table`` = device_ptr_table of shard bases.
    ``features`` bounds the valid ids (default: every row of every shard)."""
    if features is None:
        features = rows_per_shard * w_table.numel()
    _cuda(row_ptr, fid, fval, labels, w_table, bias)
    B = labels.numel()
    gs = (1.0 / B) if grad_scale is None else grad_scale
    ls = (1.0 / B) if loss_scale is None else loss_scale
    check(native.load().b200ps_sparse_lr(ptr(row_ptr), ptr(fid), ptr(fval), ptr(labels), ptr(w_table),
                                         ptr(bias), ptr(pred), ptr(dz_out), ptr(dz_in), ptr(loss_sum),
                                         ptr(global_step), B, rows_per_shard, int(features), int(forward),
                                         int(update), lr, gs, ls, stream_ptr(stream)), "sparse_lr")


def sparse_dedup_table_size(cap: int) -> int:
    """Positions of the id hash table for ``cap`` non-zeros (a power of two, at least 2 * cap)."""
    return int(native.load().b200ps_sparse_dedup_table_size(int(cap)))


def sparse_dedup_list_size(cap: int) -> int:
    """Entries of the unique list / gathered-weight / gradient buffers for ``cap`` non-zeros on the current device."""
    return int(native.load().b200ps_sparse_dedup_list_size(int(cap)))


def sparse_lr_unique(row_ptr, fid, fval, labels, w_table, bias, *, scratch, rows_per_shard, features=None, pred=None,
                     dz_out=None, dz_in=None, loss_sum=None, global_step=None, forward=True, update=True, lr=0.0,
                     grad_scale=None, loss_scale=None, unique=True, finish=True, stream=None):
    """Sparse LR step over the batch's UNIQUE ids (K4u/K5u, csrc/kernels/sparse_dedup.cu): a hash pass makes the ids
    unique (``unique=False`` reuses the list the previous launch left: the update half of a synchronous step), then one
    cooperative launch gathers each distinct weight once, walks the examples on local buffers and posts one update
    per distinct id.  ``finish`` hands the table to the next batch.  ``scratch``: SparseLR._dedup_scratch."""
    if features is None:
        features = rows_per_shard * w_table.numel()
    _cuda(row_ptr, fid, fval, labels, w_table, bias)
    B = labels.numel()
    gs = (1.0 / B) if grad_scale is None else grad_scale
    ls = (1.0 / B) if loss_scale is None else loss_scale
    sc = scratch
    check(native.load().b200ps_sparse_lr_unique(
        ptr(row_ptr), ptr(fid), ptr(fval), ptr(labels), ptr(w_table), ptr(bias), ptr(pred), ptr(dz_out), ptr(dz_in),
        ptr(loss_sum), ptr(global_step), B, fid.numel(), rows_per_shard, int(features), int(unique), int(forward),
        int(update), int(finish), lr, gs, ls, ptr(sc["keys"]), ptr(sc["slot_of"]), sc["table_size"], ptr(sc["hpos"]),
        ptr(sc["uniq"]), ptr(sc["wbuf"]), ptr(sc["gbuf"]), sc["list_cap"], ptr(sc["seg_cnt"]), ptr(sc["state"]),
        stream_ptr(stream)), "sparse_lr_unique")


def sparse_wd(row_ptr, fid, fval, labels, *, wide_table, wide_rows_per_shard, emb_table, emb_rows_per_shard, dim,
              w2_table, w2_rows_per_shard, bias, pred=None, dz_out=None, dz_in=None, loss_sum=None,
              global_step=None, forward=True, update=True, lr=0.0, grad_scale=None, loss_scale=None,
              pend_table=None, pend_rows_per_shard=0, pend2_table=None, pend2_rows_per_shard=0, features=None,
              stream=None):
    """Sparse wide & deep step over row-sharded tables (K4w/K5w): see include/b200ps.h b200ps_sparse_wd.
    With ``pend_table``/``pend2_table`` the updates are deferred (scalar coefficients; ``sparse_wd_apply``)."""
    _cuda(row_ptr, fid, fval, labels, emb_table, w2_table, bias)
    if features is None:
        features = emb_rows_per_shard * emb_table.numel()
    B = labels.numel()
    gs = (1.0 / B) if grad_scale is None else grad_scale
    ls = (1.0 / B) if loss_scale is None else loss_scale
    check(native.load().b200ps_sparse_wd(ptr(row_ptr), ptr(fid), ptr(fval), ptr(labels), ptr(wide_table),
                                         wide_rows_per_shard, ptr(emb_table), emb_rows_per_shard, dim,
                                         ptr(w2_table), w2_rows_per_shard, ptr(bias), ptr(pred), ptr(dz_out),
                                         ptr(dz_in), ptr(loss_sum), ptr(global_step), B, int(forward),
                                         int(update), lr, gs, ls, ptr(pend_table), pend_rows_per_shard,
                                         ptr(pend2_table), pend2_rows_per_shard, int(features),
                                         stream_ptr(stream)), "sparse_wd")


def sparse_wd_apply(pend_local, wide_local, emb_local, rows_local, dim, w2_table, w2_rows_per_shard, pend2_local,
                    w2_local, w2_rows_local, phase, pend2_table=None, stream=None):
    """Owner-side fold of the deferred wide & deep updates (phase 1: rows, phase 2: w2).  ``pend2_table`` (phase 1 of
    the owner-side step): also accumulate the w2 gradient ``sum_f c_f * E_old[f]`` into the pending w2 vector."""
    check(native.load().b200ps_sparse_wd_apply(ptr(pend_local), ptr(wide_local), ptr(emb_local), rows_local, dim,
                                               ptr(w2_table), w2_rows_per_shard, ptr(pend2_local), ptr(w2_local),
                                               w2_rows_local, phase, ptr(pend2_table), stream_ptr(stream)), "sparse_wd_apply")


def wd_owner_request(row_ptr, fid, *, world, rank, cap, rows_per_shard, features, inbox_table, cnt_table, resp_table,
                     count, slot_tab, stream=None):
    """Phase 1 of the owner-side wide & deep step: post this rank's row requests (include/b200ps.h)."""
    check(native.load().b200ps_wd_owner_request(ptr(row_ptr), ptr(fid), row_ptr.numel() - 1, world, rank, cap,
                                                rows_per_shard, features, ptr(inbox_table), ptr(cnt_table),
                                                ptr(resp_table), ptr(count), ptr(slot_tab), stream_ptr(stream)),
          "wd_owner_request")


def wd_owner_serve(*, world, rank, cap, inbox_table, cnt_table, resp_table, emb_local, wide_local, dim, w2_table,
                   w2_rows_per_shard, stream=None):
    """Phase 2: serve the requests in this rank's inbox from its local shards."""
    check(native.load().b200ps_wd_owner_serve(world, rank, cap, ptr(inbox_table), ptr(cnt_table), ptr(resp_table),
                                              ptr(emb_local), ptr(wide_local), dim, ptr(w2_table), w2_rows_per_shard,
                                              stream_ptr(stream)), "wd_owner_serve")


def wd_owner_fold(*, world, rank, cap, inbox_table, cnt_table, pend_local, emb_local, wide_local, dim, w2_table,
                  pend2_table, w2_rows_per_shard, stream=None):
    """Phase 4: fold the round's coefficients into this rank's rows (inbox-driven) and dw2 into the pending w2."""
    check(native.load().b200ps_wd_owner_fold(world, rank, cap, ptr(inbox_table), ptr(cnt_table), ptr(pend_local),
                                             ptr(emb_local), ptr(wide_local), dim, ptr(w2_table), ptr(pend2_table),
                                             w2_rows_per_shard, stream_ptr(stream)), "wd_owner_fold")


def wd_owner_loss(row_ptr, fid, fval, labels, *, world, rank, cap, rows_per_shard, features, resp_table, slot_tab,
                  pend_table, bias, bias_pend=None, pred=None, dz_out=None, loss_sum=None, global_step=None, update=True, lr=0.0, grad_scale=None,
                  loss_scale=None, stream=None):
    """Phase 3: loss / dz from the response array and the scalar coefficients into the owners' pending table."""
    B = labels.numel()
    gs = (1.0 / B) if grad_scale is None else grad_scale
    ls = (1.0 / B) if loss_scale is None else loss_scale
    check(native.load().b200ps_wd_owner_loss(ptr(row_ptr), ptr(fid), ptr(fval), ptr(labels), B, world, rank, cap,
                                             rows_per_shard, features, ptr(resp_table), ptr(slot_tab), ptr(pend_table), ptr(bias),
                                             ptr(bias_pend), ptr(pred), ptr(dz_out), ptr(loss_sum), ptr(global_step), int(update), lr,
                                             gs, ls, stream_ptr(stream)), "wd_owner_loss")


def auc_hist(pred, labels, thresholds, hist_pos, hist_neg, stream=None):
    _cuda(pred, labels, thresholds, hist_pos, hist_neg)
    check(native.load().b200ps_auc_hist(ptr(pred), ptr(labels), pred.numel(), ptr(thresholds),
                                        thresholds.numel(), ptr(hist_pos), ptr(hist_neg),
                                        stream_ptr(stream)), "auc_hist")


def copy_words(dst, src, n_words=None, stream=None):
    """``n_words`` 32-bit words from a device tensor into a device or pinned+mapped host tensor, as a kernel."""
    n = (src.numel() * src.element_size() // 4) if n_words is None else n_words
    check(native.load().b200ps_copy_words(ptr(dst), ptr(src), n, stream_ptr(stream)), "copy_words")


def signal_pad_words() -> int:
    return int(native.load().b200ps_signal_pad_words())


def push_sync_sgd(pads, masters, outs, grads, bounds, global_steps, *, world, rank_base, local_ranks=1,
                  max_shard_elems, dtype, replicas_to_aggregate, lr, max_ctas=0, cursor=None, cursor_step=0,
                  cursor_wrap=1, broadcast=True, stream=None):
    d = DT_BF16 if dtype == torch.bfloat16 else DT_F32
    check(native.load().b200ps_push_sync_sgd(ptr(pads), ptr(masters), ptr(outs), ptr(grads), ptr(bounds),
                                             ptr(global_steps), world, rank_base, local_ranks,
                                             max_shard_elems, d, d, replicas_to_aggregate, lr, max_ctas,
                                             ptr(cursor), cursor_step, cursor_wrap, int(broadcast),
                                             stream_ptr(stream)), "push_sync_sgd")


def push_sync_ll(pads, masters, grads, bounds, global_steps, ll_grads, ll_params, *, world, rank_base, local_ranks=1,
                 max_shard_elems, total_elems, lr, cursor=None, cursor_step=0, cursor_wrap=1, stream=None):
    """Low-latency synchronous push for small fp32 groups (see include/b200ps.h)."""
    check(native.load().b200ps_push_sync_ll(ptr(pads), ptr(masters), ptr(grads), ptr(bounds), ptr(global_steps),
                                            ptr(ll_grads), ptr(ll_params), world, rank_base, local_ranks,
                                            max_shard_elems, total_elems, lr, ptr(cursor), cursor_step, cursor_wrap,
                                            stream_ptr(stream)), "push_sync_ll")


def push_async(masters, grads, bounds, global_steps, *, world, rank_base, local_ranks=1, total_elems, lr,
               max_ctas=0, cursor=None, cursor_step=0, cursor_wrap=1, stream=None):
    check(native.load().b200ps_push_async(ptr(masters), ptr(grads), ptr(bounds), ptr(global_steps), world,
                                          rank_base, local_ranks, total_elems, lr, max_ctas, ptr(cursor),
                                          cursor_step, cursor_wrap, stream_ptr(stream)), "push_async")


def pull(masters, outs, bounds, *, world, rank_base, local_ranks=1, max_shard_elems, dtype,
         include_self=False, max_ctas=0, stream=None):
    d = DT_BF16 if dtype == torch.bfloat16 else DT_F32
    check(native.load().b200ps_pull(ptr(masters), ptr(outs), ptr(bounds), world, rank_base, local_ranks,
                                    max_shard_elems, d, int(include_self), max_ctas, stream_ptr(stream)),
          "pull")


def barrier(pads, *, world, rank_base, local_ranks=1, stream=None):
    check(native.load().b200ps_barrier(ptr(pads), world, rank_base, local_ranks, stream_ptr(stream)),
          "barrier")
