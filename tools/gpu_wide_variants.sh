#!/bin/bash
N=${1:-2}
for v in "1 1" "0 1"; do set -- $v
  B200PS_WIDE_FUSED_PULL=$1 B200PS_WIDE_OVERLAP=$2 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 297$1$2 bench.py --gpus $N --steps 10 --warmup 3 --config wide_mlp 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('fused_pull=$1 overlap=$2', round(d['ms_per_step'],3),'ms', round(d['value']))
"
done
