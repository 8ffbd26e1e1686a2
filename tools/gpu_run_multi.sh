#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L | head -8
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q -x --timeout 700 2>&1 | tail -40 > gpurun_out/test_gpu_multi.log; echo "multi exit ${PIPESTATUS[0]}"; tail -25 gpurun_out/test_gpu_multi.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29700 bench.py --gpus $N --steps 200 --warmup 20 > gpurun_out/bench_$N.json 2> gpurun_out/bench_$N.err; echo "bench exit $?"; tail -5 gpurun_out/bench_$N.err
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_$N.json').read().strip().splitlines()[-1])
    print("mnist sync", round(d["value"]), d["ms_per_step"], "e2e", round(d["e2e"]["value"]), d["n_gpus"])
    for k,v in d["extra"].items(): print(k, v.get("value"), v.get("ms_per_step"), v.get("error"))
except Exception as e: print("no bench line", e)
PY
