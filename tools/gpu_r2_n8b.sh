#!/bin/bash
mkdir -p gpurun_out
for d in 0 1; do
B200PS_WD_DEFERRED=$d timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 2956$d bench.py --gpus 8 --steps 100 --warmup 10 --config sparse_wide_deep --no-baseline 2>/dev/null | python -c "
import sys, json
d = json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('deferred=$d wide&deep N=8', round(d['value']), 'ms/step', d['ms_per_step'], 'loss', d['detail']['loss_first'], d['detail']['loss_last'])"
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29570 bench.py --gpus 8 --steps 200 --warmup 20 --config sparse_lr --no-baseline 2>/dev/null | python -c "
import sys, json
d = json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('sparse_lr N=8', round(d['value']), 'ms/step', d['ms_per_step'], d['extra'].get('sparse_lr_verify'))"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus 8 --steps 200 --warmup 20 --config mnist_mlp_sync --no-baseline 2>/dev/null | python -c "
import sys, json
d = json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('mnist sync N=8 (cooperative K13)', round(d['value']), 'us/step', d['ms_per_step']*1e3, 'e2e', round(d['e2e']['value']))"
