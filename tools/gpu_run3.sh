#!/bin/bash
mkdir -p gpurun_out
for cfg in mnist wide sparse; do
python tools/profile_step.py --config $cfg --steps 4 > gpurun_out/plain_$cfg.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_$cfg.csv python tools/profile_step.py --config $cfg --steps 4 > gpurun_out/ncu_$cfg.log 2>&1
echo "$cfg: $?"
done
