#!/bin/bash
# compute-sanitizer memcheck over the single-GPU kernel tests (one tool per call, smallest useful case)
mkdir -p gpurun_out
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 7 python -m pytest tests/test_gpu_kernels.py tests/test_zz_no_alloc.py -m gpu -q -x -k "not sparse_lr_vs" > gpurun_out/sanitizer_memcheck.log 2>&1
echo "memcheck exit $?"; tail -6 gpurun_out/sanitizer_memcheck.log
