#!/bin/bash
# 2 GPUs: multi-GPU tests (normal interpreter exit everywhere), then the driver's bench command line at N=2
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q --timeout 800 2>&1 | tail -30 > gpurun_out/pytest_multi_n2.log
echo "== pytest multi exit ${PIPESTATUS[0]}"; tail -12 gpurun_out/pytest_multi_n2.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 200 --warmup 20 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err
echo "bench exit $?"; tail -8 gpurun_out/bench_n2.err; cut -c1-1200 gpurun_out/bench_n2.json
