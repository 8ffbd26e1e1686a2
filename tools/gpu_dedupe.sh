#!/bin/bash
# unique-id sparse LR step: the kernel test, then direct vs unique on uniform and skewed ids (N ranks)
N=${1:-1}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "unique_ids or sparse_lr_vs" 2>&1 | tail -15
for skew in 1 6; do for dd in 0 1; do
  if [ "$N" = "1" ]; then
    B200PS_SPARSE_DEDUPE=$dd timeout 200 python tools/micro_dedupe.py --skew $skew 2>&1 | grep -E "sparse LR|Error|error" | tail -3
  else
    B200PS_SPARSE_DEDUPE=$dd timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2959$dd tools/micro_dedupe.py --skew $skew 2>&1 | grep -E "sparse LR|Error|error" | tail -3
  fi
done; done | tee gpurun_out/dedupe_n$N.txt
