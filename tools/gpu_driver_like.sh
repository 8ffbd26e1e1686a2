#!/bin/bash
# What the driver runs at round end, on one GPU: the GPU test-suite, smoke(), the default bench line, the reference arm.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 | tee gpurun_out/driver_like_pytest.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
timeout 600 python bench.py 2> gpurun_out/driver_like_bench.err | tee gpurun_out/driver_like_bench.json | cut -c1-600; echo "bench exit ${PIPESTATUS[0]}"
timeout 120 python bench.py --impl reference 2>/dev/null | cut -c1-300
