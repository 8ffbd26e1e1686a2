"""tcgen05/TMA GEMM vs a plain PyTorch fp32 reference (K2/K7)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(a, b, a_mn, b_mn):
    A = a.float().t() if a_mn else a.float()          # [M,K]
    Bm = b.float() if b_mn else b.float().t()          # [K,N]
    return A @ Bm


def _mk(M, K, mn, dtype, dev, g):
    shape = (K, M) if mn else (M, K)
    return (torch.randn(shape, device=dev, generator=g) * 0.5).to(dtype)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K", [(128, 256, 128), (256, 512, 320), (100, 100, 784), (784, 100, 100),
                                   (300, 200, 72), (128, 64, 64), (1000, 1000, 1000)])
def test_gemm_layouts(native_lib, dtype, a_mn, b_mn, M, N, K):
    from b200ps import ops
    dev = "cuda"
    esz = 2 if dtype == torch.bfloat16 else 4
    if ((M if a_mn else K) * esz) % 16 or ((N if b_mn else K) * esz) % 16:
        pytest.skip("row stride not 16-byte aligned")
    g = torch.Generator(device=dev).manual_seed(M * 7 + N * 3 + K)
    a = _mk(M, K, a_mn, dtype, dev, g)
    b = _mk(N, K, b_mn, dtype, dev, g)
    out = ops.gemm(a, b, a_mn_major=a_mn, b_mn_major=b_mn, out_dtype=torch.float32)
    torch.cuda.synchronize()
    ref = _ref(a, b, a_mn, b_mn)
    tol = 2e-2 if dtype == torch.bfloat16 else 5e-3
    err = (out - ref).abs().max().item()
    scale = ref.abs().max().item() + 1e-6
    assert err / scale < tol, f"max err {err} (scale {scale})"


def test_gemm_bias_sigmoid_bf16_out(native_lib):
    from b200ps import ops
    g = torch.Generator(device="cuda").manual_seed(0)
    M, N, K = 384, 512, 256
    a = (torch.randn(M, K, device="cuda", generator=g) * 0.3).bfloat16()
    w = (torch.randn(K, N, device="cuda", generator=g) * 0.3).bfloat16()   # [in,out] => MN-major B
    bias = torch.randn(N, device="cuda", generator=g)
    out = ops.gemm(a, w, b_mn_major=True, bias=bias, act=ops.ACT_SIGMOID)
    ref = torch.sigmoid(a.float() @ w.float() + bias)
    assert out.dtype == torch.bfloat16
    assert (out.float() - ref).abs().max().item() < 1.5e-2


def test_gemm_dsigmoid_epilogue(native_lib):
    from b200ps import ops
    g = torch.Generator(device="cuda").manual_seed(1)
    M, N, K = 256, 384, 512
    dz = (torch.randn(M, K, device="cuda", generator=g) * 0.1).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) * 0.1).bfloat16()   # [in=N,out=K]: K-major B
    act = torch.rand(M, N, device="cuda", generator=g).bfloat16()
    out = ops.gemm(dz, w, dsigmoid_of=act)
    ref = (dz.float() @ w.float().t()) * act.float() * (1 - act.float())
    assert (out.float() - ref).abs().max().item() < 2e-2 * ref.abs().max().item() + 1e-3


def test_gemm_fused_colsum(native_lib):
    from b200ps import ops
    g = torch.Generator(device="cuda").manual_seed(5)
    M, N, K = 512, 320, 256
    dz = (torch.randn(M, K, device="cuda", generator=g) * 0.1).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) * 0.1).bfloat16()
    act = torch.rand(M, N, device="cuda", generator=g).bfloat16()
    cs = torch.full((N,), 123.0, device="cuda")
    out = ops.gemm(dz, w, dsigmoid_of=act, colsum_out=cs)
    ref = (dz.float() @ w.float().t()) * act.float() * (1 - act.float())
    assert (out.float() - ref).abs().max().item() < 2e-2 * ref.abs().max().item() + 1e-3
    assert torch.allclose(cs, ref.sum(0), atol=2e-2 * ref.abs().sum(0).max().item())


def test_gemm_splitk_atomic(native_lib):
    from b200ps import ops
    g = torch.Generator(device="cuda").manual_seed(2)
    M, N, K = 100, 100, 784
    x = torch.rand(M, K, device="cuda", generator=g)
    w = torch.randn(K, N, device="cuda", generator=g)
    out = torch.zeros(M, N, device="cuda")
    ops.gemm(x, w, out, b_mn_major=True, epilogue=ops.EPI_ATOMIC_ADD, split_k=5)
    ref = x @ w
    assert (out - ref).abs().max().item() < 5e-3 * ref.abs().max().item()


def test_gemm_fused_sgd(native_lib):
    from b200ps import ops
    g = torch.Generator(device="cuda").manual_seed(3)
    Bn, I, O = 256, 512, 384
    x = (torch.randn(Bn, I, device="cuda", generator=g) * 0.3).bfloat16()
    dz = (torch.randn(Bn, O, device="cuda", generator=g) * 0.3).bfloat16()
    master = torch.randn(I, O, device="cuda", generator=g)
    lowp = master.bfloat16()
    want = master - 0.1 * (x.float().t() @ dz.float())
    ops.gemm(x, dz, a_mn_major=True, b_mn_major=True, epilogue=ops.EPI_SGD, sgd_master=master,
             sgd_lowp=lowp, sgd_lr=0.1)
    assert (master - want).abs().max().item() < 2e-2 * want.abs().max().item()
    assert torch.equal(lowp, master.bfloat16())


def test_gemm_rejects_bad_args(native_lib):
    from b200ps import ops
    a = torch.zeros(10, 10, device="cuda")
    with pytest.raises(ops.native.NativeError):
        ops.gemm(a[:, :7].contiguous(), a[:, :7].contiguous())  # 28-byte rows: not TMA-able


@pytest.mark.parametrize("world,rank", [(1, 0), (2, 1), (4, 0), (4, 2), (8, 5)])
def test_fused_pull_gemm_virtual_ranks(native_lib, world, rank):
    """K2: the weight all-gather runs inside the GEMM launch (gather CTAs + GEMM CTAs)."""
    from b200ps import ops
    g = torch.Generator(device="cuda").manual_seed(world * 10 + rank)
    M, K, N = 384, 1024, 512
    x = (torch.randn(M, K, device="cuda", generator=g) * 0.3).bfloat16()
    w_full = (torch.randn(K, N, device="cuda", generator=g) * 0.3).bfloat16()
    rows = K // world
    # every "rank" holds a full-size buffer that is only valid in its own row range
    copies = []
    for r in range(world):
        c = torch.full((K, N), float("nan"), device="cuda", dtype=torch.bfloat16)
        c[r * rows:(r + 1) * rows] = w_full[r * rows:(r + 1) * rows]
        copies.append(c)
    table = ops.device_ptr_table([c.data_ptr() for c in copies], "cuda")
    flags = torch.zeros(K // 64 + 1, device="cuda", dtype=torch.int32)
    bias = torch.randn(N, device="cuda", generator=g)
    ref = torch.sigmoid(x.float() @ w_full.float() + bias)
    for _ in range(2):                                   # second launch: flags were re-armed by the first
        out = ops.gemm(x, copies[rank], b_mn_major=True, bias=bias, act=ops.ACT_SIGMOID,
                       gather=dict(src_table=table, flags=flags, world=world, rank=rank, ctas=3))
        torch.cuda.synchronize()
        assert (out.float() - ref).abs().max().item() < 1.5e-2
        assert torch.equal(copies[rank], w_full)         # the all-gather happened as a side effect
        assert int(flags.abs().sum().item()) == 0
        if world > 1:                                     # invalidate the gathered part again
            keep = copies[rank][rank * rows:(rank + 1) * rows].clone()
            copies[rank].fill_(float("nan")); copies[rank][rank * rows:(rank + 1) * rows] = keep
