#!/bin/bash
mkdir -p gpurun_out
for f in tests/test_gpu_gemm.py tests/test_gpu_kernels.py tests/test_gpu_train.py; do
  timeout 600 python -m pytest $f -m gpu -q --maxfail 12 --timeout 300 2>&1 | tail -80 > gpurun_out/$(basename $f .py).log
  echo "== $f exit ${PIPESTATUS[0]}"; grep -E "^(FAILED|ERROR|[0-9]+ (passed|failed))|passed|failed" gpurun_out/$(basename $f .py).log | tail -15
done
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -5
timeout 600 python bench.py --steps 200 --warmup 20 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; tail -3 gpurun_out/bench.err; cat gpurun_out/bench.json
