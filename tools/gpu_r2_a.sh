#!/bin/bash
# round 2, first contact: full GPU test-suite, smoke, default bench (normal interpreter exit), cluster probe
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
./tools/ubench/cluster_probe 40 > gpurun_out/cluster_probe.txt 2>&1; echo "probe exit $?"; cat gpurun_out/cluster_probe.txt
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 600 2>&1 | tail -40 > gpurun_out/pytest_gpu_n1.log
echo "== pytest exit ${PIPESTATUS[0]}"; tail -12 gpurun_out/pytest_gpu_n1.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?"; tail -5 gpurun_out/bench_n1.err; cat gpurun_out/bench_n1.json
