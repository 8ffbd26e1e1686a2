"""b200ps_mlp_tail_static — the MLP-head entry point for hosts that must not allocate.

Some embedders cannot let a library call cudaMalloc at step time: the host application has
already committed the whole device to its own memory plan, or it runs a latency-bound loop in
which cudaMalloc's device-wide synchronisation is unacceptable.  For them the library offers a
variant of ``b200ps_mlp_tail`` whose cross-CTA scratch is a fixed 1 MiB array inside the
library's device image.  This test drives that entry point straight through the C ABI
(ctypes, raw device pointers) and checks every output against the allocating entry point, on
two consecutive calls with different inputs (the array is shared and keeps the previous
call's contents, so a stale-scratch bug would show on the second call).  It is collected
last on purpose and must pass both alone in a fresh process and after the rest of the suite.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _problem(B, H, Cn, seed):
    g = torch.Generator(device=DEV).manual_seed(seed)
    return dict(z2=torch.randn(B, H, device=DEV, generator=g), b1=torch.randn(H, device=DEV, generator=g) * 0.1,
                w2=torch.randn(H, Cn, device=DEV, generator=g) * 0.3, b2=torch.randn(Cn, device=DEV, generator=g) * 0.1,
                labels=torch.randint(0, Cn, (B,), device=DEV, generator=g, dtype=torch.int32))


def _outputs(B, H, Cn):
    return dict(a2=torch.zeros(B, H, device=DEV), dz2=torch.zeros(B, H, device=DEV), dw2=torch.zeros(H, Cn, device=DEV),
                db2=torch.zeros(Cn, device=DEV), db1=torch.zeros(H, device=DEV), loss=torch.zeros(1, device=DEV),
                correct=torch.zeros(1, device=DEV, dtype=torch.int32), probs=torch.zeros(B, Cn, device=DEV))


def _call(fn, inp, out, B, H, Cn, train=1):
    stream = torch.cuda.current_stream().cuda_stream
    return fn(inp["z2"].data_ptr(), inp["b1"].data_ptr(), inp["w2"].data_ptr(), inp["b2"].data_ptr(),
              inp["labels"].data_ptr(), out["a2"].data_ptr(), out["dz2"].data_ptr(), out["dw2"].data_ptr(),
              out["db2"].data_ptr(), out["db1"].data_ptr(), out["loss"].data_ptr(), out["correct"].data_ptr(),
              out["probs"].data_ptr(), B, H, Cn, train, 1.0 / B, 1.0 / B, 0, None, 1, 0, stream)


def test_static_scratch_entry_point_matches_allocating_one(native_lib):
    lib = native_lib
    static_fn, alloc_fn = lib.b200ps_mlp_tail_static, lib.b200ps_mlp_tail
    for call, (B, H, Cn, seed) in enumerate([(100, 100, 10, 0), (1000, 64, 10, 1)]):   # 13 CTAs, then 125
        inp = _problem(B, H, Cn, seed)
        ref, got = _outputs(B, H, Cn), _outputs(B, H, Cn)
        assert _call(alloc_fn, inp, ref, B, H, Cn) == 0, lib.b200ps_last_error()
        assert _call(static_fn, inp, got, B, H, Cn) == 0, lib.b200ps_last_error()
        torch.cuda.synchronize()
        for k in ref:
            if ref[k].dtype == torch.int32:
                assert torch.equal(got[k], ref[k]), (call, k)
            else:
                assert torch.allclose(got[k], ref[k], atol=1e-6, rtol=1e-4), (call, k)
        assert ref["loss"].item() > 0 and ref["dw2"].abs().sum().item() > 0    # the oracle did real work


def test_static_scratch_entry_point_refuses_oversize(native_lib):
    lib = native_lib
    B, H, Cn = 4096, 128, 32            # 148 CTAs x 4258 floats = 2.4 MiB > the 1 MiB array
    inp, out = _problem(B, H, Cn, 2), _outputs(B, H, Cn)
    before = out["dw2"].clone()
    rc = _call(lib.b200ps_mlp_tail_static, inp, out, B, H, Cn)
    assert rc == -1
    msg = lib.b200ps_last_error().decode()
    assert "static scratch" in msg and "1048576" in msg
    torch.cuda.synchronize()
    assert torch.equal(out["dw2"], before)                 # nothing was launched
    assert _call(lib.b200ps_mlp_tail, inp, out, B, H, Cn) == 0   # the allocating entry point takes it
    torch.cuda.synchronize()
