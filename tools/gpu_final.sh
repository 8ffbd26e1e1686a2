#!/bin/bash
# Round-end measurement: both arms, every config, on N GPUs -> gpurun_out/final_{b200ps,baseline}_nN.json
N=${1:-1}
mkdir -p gpurun_out
if [ "$N" = "1" ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2972$N"; fi
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv -lms 500 > gpurun_out/clocks_n$N.csv 2>/dev/null &
SMI=$!
timeout 1200 $L bench.py --gpus $N --steps 200 --warmup 20 2> gpurun_out/final_b200ps_n$N.err | grep '^{' > gpurun_out/final_b200ps_n$N.json; echo "b200ps exit ${PIPESTATUS[0]}"
timeout 1200 $L bench.py --impl baseline --gpus $N --steps 200 --warmup 20 2> gpurun_out/final_baseline_n$N.err | grep '^{' > gpurun_out/final_baseline_n$N.json; echo "baseline exit ${PIPESTATUS[0]}"
kill $SMI 2>/dev/null
python - <<PY
import json
for name in ("final_b200ps_n$N", "final_baseline_n$N"):
    try:
        d = json.loads(open('gpurun_out/%s.json' % name).read().strip().splitlines()[-1])
        print(name, round(d["value"]), round(d["ms_per_step"] * 1e3, 2), "us e2e", round((d.get("e2e") or {}).get("value", 0)), "vs_baseline", d.get("vs_baseline"), (d.get("graphed") or {}).get("value"))
        for k, v in (d.get("extra") or {}).items():
            print("   ", k, round(v.get("value", 0)) if "value" in v else v, v.get("ms_per_step"), (v.get("e2e") or {}).get("value"), (v.get("graphed") or {}).get("value"))
    except Exception as e:
        print(name, "no line", e)
PY
