#!/bin/bash
# 2 GPUs: owner-side wide & deep step against the direct kernel
mkdir -p gpurun_out
for o in 2 4; do
B200PS_WD_OWNER_MIN=$o timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2958$o bench.py --gpus 2 --steps 100 --warmup 10 --config sparse_wide_deep --no-baseline 2>gpurun_out/wd2_$o.err | tee gpurun_out/wd2_ownermin$o.json | python -c "
import sys, json
d = json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('owner_min=$o wide&deep N=2', round(d['value']), 'ms/step', d['ms_per_step'], 'e2e', round(d['e2e']['value']), d['config']['mode'], 'loss', d['detail']['loss_first'], d['detail']['loss_last'])" || tail -5 gpurun_out/wd2_$o.err
done
