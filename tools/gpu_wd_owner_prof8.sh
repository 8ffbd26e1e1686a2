#!/bin/bash
# owner-side wide & deep on 8 VIRTUAL ranks of one GPU: plain run, then ONE ncu invocation
#   mode "list" (default): per-kernel durations of one step (48 launches)   mode "full": --set full of serve and fold
mkdir -p gpurun_out
export WD_WORLD=8
python tools/prof_wd_owner.py > gpurun_out/plain_wd_owner8.log 2>&1 || { tail -5 gpurun_out/plain_wd_owner8.log; exit 1; }
if [ "${1:-list}" = "list" ]; then
  ncu --metrics gpu__time_duration.sum,dram__bytes.sum,launch__grid_size --clock-control none -k regex:'wd_' -s 100 -c 48 --csv --log-file gpurun_out/wd_owner8_launches.csv python tools/prof_wd_owner.py > gpurun_out/ncu_wd_owner8a.log 2>&1
else
  ncu --set full --clock-control none --import-source on -k regex:'wd_serve|wd_fold' -s 24 -c 4 -o gpurun_out/prof_wd_owner8 -f python tools/prof_wd_owner.py > gpurun_out/ncu_wd_owner8.log 2>&1
  tail -2 gpurun_out/ncu_wd_owner8.log
fi
