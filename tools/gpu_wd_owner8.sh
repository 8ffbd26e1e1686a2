#!/bin/bash
# 8-GPU check of the owner-side wide & deep step: oracle comparison, then the bench config (owner-side only).
N=${1:-8}
mkdir -p gpurun_out
python -m pytest tests/test_gpu_multi.py -m gpu -x -q -k "oracles" 2>&1 | tail -3
B200PS_WD_OWNER=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29581 bench.py --gpus $N --steps 100 --warmup 10 --config sparse_wide_deep --no-baseline 2>gpurun_out/wd_1.err | tee gpurun_out/wd_owner_n$N.json | python -c "
import sys, json
d = json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('owner=1 wide&deep N=$N', round(d['value']), 'ms/step', d['ms_per_step'], 'e2e', round(d['e2e']['value']), d['config']['mode'], 'loss', d['detail']['loss_first'], d['detail']['loss_last'])" || tail -5 gpurun_out/wd_1.err
PYTHONPATH=. timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29583 tools/micro_wd_owner.py 2>&1 | grep -v Warning | tail -12
