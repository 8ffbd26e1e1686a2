#!/bin/bash
N=$1
for o in 1 0; do
B200PS_WD_OWNER=$o timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2958$o bench.py --gpus $N --steps 100 --warmup 10 --config sparse_wide_deep --no-baseline 2>gpurun_out/wd_$o.err | python -c "
import sys, json
d = json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('owner=$o wide&deep N=$N', round(d['value']), 'ms/step', d['ms_per_step'], 'e2e', round(d['e2e']['value']), d['config']['mode'], 'loss', d['detail']['loss_first'], d['detail']['loss_last'])" || tail -5 gpurun_out/wd_$o.err
done
