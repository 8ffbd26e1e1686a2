"""BASELINE.json config 4 (8 x 8192 x 8192 bf16 MLP, 4096 examples/GPU, sync PS) — a thin front end of ../bench.py.

    python bench/wide_mlp.py --steps 200 --warmup 20            # 1 GPU
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 bench/wide_mlp.py --gpus 8
"""
import os
import runpy
import sys

if not any(a.startswith("--config") for a in sys.argv[1:]):
    sys.argv += ["--config", "wide_mlp"]
sys.argv[0] = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "bench.py")
runpy.run_path(sys.argv[0], run_name="__main__")
