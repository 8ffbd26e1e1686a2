"""8-GPU experiment: the wide bf16 MLP step (BASELINE.json config 4) with the per-layer pushes on the P2P kernel vs
through the switch (K1-NVLS) at several CTA caps.  torchrun, one rank per GPU."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ps import ops  # noqa: E402
from b200ps.parallel.context import init_from_env  # noqa: E402
from b200ps.train import WideMLPTrainer  # noqa: E402

ctx = init_from_env()
dev = ctx.device
B, L, D = 4096, 8, 8192
xs = torch.empty(2, B, D, device=dev, dtype=torch.bfloat16); ys = torch.empty(2, B, device=dev, dtype=torch.int32)
for i in range(2):
    ops.synth_dense_bf16(xs[i], ys[i], classes=D, seed=3, offset=(ctx.rank * 2 + i) * B)
configs = [("p2p", None)] + [("nvls", c) for c in (16, 32, 64, 148)]
if len(sys.argv) > 1:
    configs = [("nvls", int(c)) if c != "p2p" else ("p2p", None) for c in sys.argv[1].split(",")]
for path, cap in configs:
    os.environ.pop("B200PS_NVLS_CTAS", None)
    if cap:
        os.environ["B200PS_NVLS_CTAS"] = str(cap)
    tr = WideMLPTrainer(ctx, batch_size=B, layers=L, hidden=D, lr=0.01, use_graph=True)
    tr.store.use_nvls = path == "nvls"
    for i in range(5):
        tr.step(xs[i % 2], ys[i % 2])
    ctx.control_barrier(); torch.cuda.synchronize()
    st, en = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    st.record()
    for i in range(10):
        tr.step(xs[i % 2], ys[i % 2])
    en.record(); torch.cuda.synchronize()
    ms = ctx.all_reduce_max(st.elapsed_time(en) / 10)
    if ctx.rank == 0:
        print(json.dumps({"push": path, "nvls_cta_cap": cap, "ms_per_step": round(ms, 3), "loss": tr.read_loss(),
                          "nvls_memory": bool(tr.store.groups["bf16:0"].work.mc_ptr)}), flush=True)
    else:
        tr.read_loss()
    del tr
    import gc; gc.collect(); torch.cuda.empty_cache()
    ctx.control_barrier()
ctx.shutdown()
