#!/bin/bash
# round-end style evaluation on an 8-GPU box: tests, both bench arms at N = 1, 2, 4, 8, push/pull sweep at N = 8
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 800 2>&1 | tail -3
for N in 1 2 4 8; do bash tools/gpu_baseline.sh $N 2>&1 | grep -A4 "^bench_\|^baseline_"; done
SWEEP_MAX_BYTES=1073741824 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29790 bench/pushpull_sweep.py 2>&1 | grep "^{" | tail -3
