#!/bin/bash
# 2 GPUs: direct vs unique-id sparse LR step on skewed ids (half of the hot ids live on the other GPU)
mkdir -p gpurun_out
for skew in 6 12; do for dd in 0 1; do
  B200PS_SPARSE_DEDUPE=$dd timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2959$dd tools/micro_dedupe.py --skew $skew 2>&1 | grep -E "sparse LR|Error" | tail -2
done; done | tee gpurun_out/dedupe_n2.txt
