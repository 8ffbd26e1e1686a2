#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q --timeout 800 -k "oracles or trainers" 2>&1 | tail -40 > gpurun_out/pytest_multi3_n2.log
echo "== pytest exit ${PIPESTATUS[0]}"; tail -30 gpurun_out/pytest_multi3_n2.log
grep "^rank 0" gpurun_out/multi_gpu_oracle_worker_n2.log | cut -c1-1800
timeout 300 python -m pytest tests/test_io.py -m gpu -x -q 2>&1 | tail -5
