#!/bin/bash
mkdir -p gpurun_out
python tools/profile_step.py --config wide --steps 2 > gpurun_out/plain_wide.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'gemm_tcgen05' -s 30 -c 3 -o gpurun_out/prof_wide -f python tools/profile_step.py --config wide --steps 2 > gpurun_out/ncu_wide.log 2>&1
echo "ncu exit $?"; tail -2 gpurun_out/ncu_wide.log
