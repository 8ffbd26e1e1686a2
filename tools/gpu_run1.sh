#!/bin/bash
# first GPU contact: kernel numerics, then a GEMM micro-benchmark
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
for f in tests/test_gpu_gemm.py tests/test_gpu_kernels.py; do
  timeout 600 python -m pytest $f -m gpu -q --maxfail 12 --timeout 300 2>&1 | tail -60 > gpurun_out/$(basename $f .py).log
  echo "== $f exit ${PIPESTATUS[0]}"; tail -15 gpurun_out/$(basename $f .py).log
done
timeout 300 python bench/gemm_bench.py > gpurun_out/gemm_bench.log 2>&1; echo "gemm_bench exit $?"; tail -20 gpurun_out/gemm_bench.log
