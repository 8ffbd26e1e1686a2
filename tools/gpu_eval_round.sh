#!/bin/bash
# b200ps arm (all configs) + the baseline arm for one config on N GPUs; writes gpurun_out/{bench,baselinewd}_N.json
N=${1:-2}; BCFG=${2:-sparse_wide_deep}
mkdir -p gpurun_out
if [ "$N" = "1" ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2971$N"; fi
timeout 900 $L bench.py --gpus $N --steps 200 --warmup 20 > gpurun_out/bench_$N.json 2> gpurun_out/bench_$N.err; echo "bench exit $?"
timeout 600 $L bench.py --impl baseline --config $BCFG --gpus $N --steps 200 --warmup 20 > gpurun_out/baselinewd_$N.json 2> gpurun_out/baselinewd_$N.err; echo "baseline($BCFG) exit $?"
python - <<PY
import json
for name in ("bench_$N","baselinewd_$N"):
    try:
        d=[json.loads(l) for l in open('gpurun_out/%s.json'%name) if l.startswith('{')][-1]
        print(name, d.get("config",{}).get("model"), round(d["value"]), round(d["ms_per_step"]*1e3,1),"us", "e2e", round((d.get("e2e") or {}).get("value",0)))
        for k,v in (d.get("extra") or {}).items(): print("   ", k, round(v.get("value",0)), v.get("ms_per_step"), v.get("error"))
    except Exception as e: print(name, "no line", e)
PY
