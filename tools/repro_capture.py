import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ps import ops
from b200ps.parallel.context import init_from_env
from b200ps.train import MLPTrainer
mode = sys.argv[1]
if mode in ("bf16", "k2"):
    x = torch.randn(256, 512, device="cuda").bfloat16(); w = torch.randn(512, 512, device="cuda").bfloat16()
    if mode == "k2":
        table = ops.device_ptr_table([w.data_ptr()] * 2, "cuda"); flags = torch.zeros(512 // 64 + 1, device="cuda", dtype=torch.int32)
        ops.gemm(x, w, b_mn_major=True, gather=dict(src_table=table, flags=flags, world=2, rank=0))
    else:
        ops.gemm(x, w, b_mn_major=True)
    torch.cuda.synchronize()
if mode == "f32big":
    x = torch.randn(256, 512, device="cuda"); w = torch.randn(512, 512, device="cuda")
    ops.gemm(x, w, b_mn_major=True); torch.cuda.synchronize()
xs = torch.empty(700, 784, device="cuda"); ys = torch.empty(700, device="cuda", dtype=torch.int32)
ops.synth_mnist(xs, ys, seed=1)
a = MLPTrainer(init_from_env(virtual_world=1), batch_size=100, lr=0.01, mode="sync", use_graph=True)
a.attach_dataset(xs, ys)
try:
    a.run_resident(11)
    torch.cuda.synchronize()
    print(mode, "OK")
except Exception as e:
    print(mode, "FAIL", str(e)[:150])
