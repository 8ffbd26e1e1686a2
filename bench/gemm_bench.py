"""Micro-benchmark of the tcgen05 GEMM against torch.matmul (cuBLAS) on the wide-MLP shapes."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ps import ops  # noqa: E402


def timeit(fn, iters=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    st, en = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    st.record()
    for _ in range(iters):
        fn()
    en.record()
    torch.cuda.synchronize()
    return st.elapsed_time(en) / iters


def main():
    dev = "cuda"
    out = []
    for (M, N, K) in [(4096, 8192, 8192), (8192, 8192, 4096), (2048, 2048, 2048), (1024, 8192, 8192)]:
        for name, a_mn, b_mn in [("fwd  A[M,K] B[K,N]", False, True), ("dgrad A[M,K] B[N,K]", False, False),
                                 ("wgrad A[K,M] B[K,N]", True, True)]:
            a = torch.randn((K, M) if a_mn else (M, K), device=dev).bfloat16()
            b = torch.randn((K, N) if b_mn else (N, K), device=dev).bfloat16()
            c = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
            A = a.t() if a_mn else a
            Bm = b if b_mn else b.t()
            t_mine = timeit(lambda: ops.gemm(a, b, c, a_mn_major=a_mn, b_mn_major=b_mn))
            t_ref = timeit(lambda: torch.matmul(A, Bm, out=c))
            fl = 2.0 * M * N * K
            rec = dict(shape=[M, N, K], layout=name, b200ps_ms=round(t_mine, 4), cublas_ms=round(t_ref, 4),
                       b200ps_tflops=round(fl / t_mine / 1e9, 1), cublas_tflops=round(fl / t_ref / 1e9, 1))
            print(json.dumps(rec), flush=True)
            out.append(rec)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open("gpurun_out/gemm_bench.json", "w"), indent=1)


if __name__ == "__main__":
    main()
