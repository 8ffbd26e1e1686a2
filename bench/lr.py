"""BASELINE.json config 3 (sparse LR, 2^24 hashed features, ~100 nnz/example) — a thin front end of ../bench.py.

    python bench/lr.py --steps 200 --warmup 20            # 1 GPU
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 bench/lr.py --gpus 8
"""
import os
import runpy
import sys

if not any(a.startswith("--config") for a in sys.argv[1:]):
    sys.argv += ["--config", "sparse_lr"]
sys.argv[0] = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "bench.py")
runpy.run_path(sys.argv[0], run_name="__main__")
