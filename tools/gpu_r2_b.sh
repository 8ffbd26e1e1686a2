#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q --timeout 300 -k "cluster or fused or sparse_kernels_refuse" 2>&1 | tail -40 > gpurun_out/pytest_b.log
echo "== pytest exit ${PIPESTATUS[0]}"; tail -30 gpurun_out/pytest_b.log
timeout 900 python -m pytest tests/test_gpu_train.py -m gpu -x -q --timeout 300 2>&1 | tail -30 > gpurun_out/pytest_b2.log
echo "== pytest train exit ${PIPESTATUS[0]}"; tail -20 gpurun_out/pytest_b2.log
timeout 600 python bench.py --config mnist_mlp_sync --no-baseline > gpurun_out/bench_b.json 2> gpurun_out/bench_b.err; echo "bench exit $?"; tail -5 gpurun_out/bench_b.err; cut -c1-1500 gpurun_out/bench_b.json
