#!/bin/bash
# measure the PyTorch+NCCL baseline and the b200ps arm back to back on N GPUs
N=${1:-1}
mkdir -p gpurun_out
if [ "$N" = "1" ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29701"; fi
timeout 900 $L bench.py --impl baseline --gpus $N --steps 200 --warmup 20 > gpurun_out/baseline_$N.json 2> gpurun_out/baseline_$N.err; echo "baseline exit $?"; tail -3 gpurun_out/baseline_$N.err
timeout 900 $L bench.py --gpus $N --steps 200 --warmup 20 > gpurun_out/bench_$N.json 2> gpurun_out/bench_$N.err; echo "bench exit $?"; tail -3 gpurun_out/bench_$N.err
python - <<PY
import json
for name in ("baseline_$N","bench_$N"):
    try:
        d=json.loads(open('gpurun_out/%s.json'%name).read().strip().splitlines()[-1])
        print(name, "mnist_sync", round(d["value"]), round(d["ms_per_step"]*1e3,1),"us", "e2e", round((d.get("e2e") or {}).get("value",0)))
        for k,v in d["extra"].items(): print("   ", k, round(v.get("value",0)), v.get("ms_per_step"), v.get("error"))
    except Exception as e: print(name, "no line", e)
PY
