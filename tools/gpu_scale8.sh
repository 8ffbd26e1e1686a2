#!/bin/bash
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q -x --timeout 700 2>&1 | tail -4
for N in 2 4 8; do bash tools/gpu_baseline.sh $N; done
