#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q --timeout 120 -k "cluster" 2>&1 | tail -30 > gpurun_out/pytest_c.log
echo "== pytest exit ${PIPESTATUS[0]}"; tail -25 gpurun_out/pytest_c.log
PYTHONPATH=. timeout 120 python tools/micro_cluster.py 2>&1 | tail -12
