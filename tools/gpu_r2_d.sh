#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_train.py -m gpu -x -q --timeout 300 -k "sparse" 2>&1 | tail -8
timeout 600 python bench.py --config sparse_lr --no-baseline > gpurun_out/bench_sparse.json 2> gpurun_out/bench_sparse.err; echo "bench exit $?"; tail -3 gpurun_out/bench_sparse.err
python -c "
import json
d = json.loads([l for l in open('gpurun_out/bench_sparse.json') if l.startswith('{')][-1])
print('sparse_lr', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value']); print(d['extra'])
"
