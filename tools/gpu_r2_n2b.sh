#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_multi.py -m gpu -x -q --timeout 800 -k "oracles or dead_rank or skew" 2>&1 | tail -40 > gpurun_out/pytest_multi2_n2.log
echo "== pytest exit ${PIPESTATUS[0]}"; tail -30 gpurun_out/pytest_multi2_n2.log
for f in gpurun_out/multi_gpu_*_n2.log; do echo "--- $f"; grep "^rank 0" $f | cut -c1-900; done
