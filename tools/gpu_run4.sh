#!/bin/bash
mkdir -p gpurun_out
for f in tests/test_gpu_kernels.py tests/test_gpu_train.py; do
  timeout 600 python -m pytest $f -m gpu -q --maxfail 12 --timeout 300 2>&1 | tail -80 > gpurun_out/$(basename $f .py).log
  echo "== $f exit ${PIPESTATUS[0]}"; grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/$(basename $f .py).log | tail -15
done
python tools/micro_mnist.py 2>&1 | grep -E "us$|rror"
timeout 600 python bench.py --steps 200 --warmup 20 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; tail -3 gpurun_out/bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench.json').read().strip().splitlines()[-1])
print("mnist sync", round(d["value"]), d["ms_per_step"], "e2e", round(d["e2e"]["value"]), d["clocks"])
for k,v in d["extra"].items(): print(k, v.get("value"), v.get("ms_per_step"), v.get("e2e",{}).get("value"), v.get("error"))
PY
