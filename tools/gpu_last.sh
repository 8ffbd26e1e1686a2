#!/bin/bash
# last call of the round: ncu --set full of the unique-id step (after a plain run of the same command), then the driver's sequence
mkdir -p gpurun_out
B200PS_SPARSE_DEDUPE=1 timeout 120 python tools/micro_dedupe.py --skew 6 2>&1 | grep "sparse LR" || exit 1
B200PS_SPARSE_DEDUPE=1 timeout 300 ncu --set full --clock-control none --import-source on -k regex:sparse_lr_unique -s 60 -c 1 -o gpurun_out/prof_dedupe -f python tools/micro_dedupe.py --skew 6 > gpurun_out/ncu_dedupe_full.log 2>&1
tail -2 gpurun_out/ncu_dedupe_full.log
bash tools/gpu_driver_like.sh
