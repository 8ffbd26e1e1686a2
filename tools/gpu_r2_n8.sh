#!/bin/bash
# 8 GPUs: push/pull sweep (P2P vs NVLS vs NCCL), the driver's bench line, the multi-GPU test-suite
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench/pushpull_sweep.py > gpurun_out/sweep_n8.log 2>&1
echo "sweep exit $?"; grep '^{' gpurun_out/sweep_n8.log | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    print(d['bytes'], d['push_sync_path'], 'push', d['push_sync_us'], 'p2p', d['push_sync_p2p_us'], 'nvls', d['push_sync_nvls_us'], 'nccl', d['nccl_push_us'], '| pull', d['pull_us'], 'nvls', d['pull_nvls_us'], 'nccl', d['nccl_allgather_us'], '| async', d['push_async_us'], 'nccl', d['nccl_rs_apply_us'])
"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 8 --steps 200 --warmup 20 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err
echo "bench exit $?"; tail -3 gpurun_out/bench_n8.err; python -c "
import json
d = json.loads([l for l in open('gpurun_out/bench_n8.json') if l.startswith('{')][-1])
print('N=8 value', d['value'], 'ms/step', d['ms_per_step'], 'e2e', d['e2e']['value'], 'vs_baseline', d['vs_baseline'])
for k, v in d['extra'].items():
    print(' ', k, v.get('value'), v.get('ms_per_step'), v.get('error'))
"
timeout 1200 python -m pytest tests/test_gpu_multi.py -m gpu -x -q --timeout 900 2>&1 | tail -15 > gpurun_out/pytest_multi_n8.log
echo "== pytest exit ${PIPESTATUS[0]}"; tail -8 gpurun_out/pytest_multi_n8.log
for f in gpurun_out/multi_gpu_*_n8.log; do echo "--- $f"; grep "^rank 0" $f | cut -c1-1500; done
