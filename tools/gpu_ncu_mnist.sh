#!/bin/bash
mkdir -p gpurun_out
python tools/profile_step.py --config mnist --steps 6 > gpurun_out/plain_mnist.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'mlp_tail|push_sync|gemm_tcgen05' -s 16 -c 4 -o gpurun_out/prof_mnist -f python tools/profile_step.py --config mnist --steps 6 > gpurun_out/ncu_mnist.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/ncu_mnist.log; ls -la gpurun_out/*.ncu-rep
