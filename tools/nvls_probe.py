"""Probe (2+ GPUs, torchrun): is NVLS multicast usable on this box, through torch's symmetric memory?

Prints, per rank, the peer pointers and the multicast pointer of a symmetric allocation, runs torch's own
multimem all-reduce on it, and reports what failed if anything did.  Input for the NVLS (`multimem.ld_reduce` /
`multimem.st`) variant of the push kernels planned next: the b200ps pool allocates with cudaMalloc + CUDA IPC, which
cannot be bound to a multicast object; this checks whether torch's allocator can stand in for it.
"""
import os
import traceback

import torch
import torch.distributed as dist


def main():
    dist.init_process_group("nccl")
    rank, world = dist.get_rank(), dist.get_world_size()
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    out = {"rank": rank, "world": world, "torch": torch.__version__}
    try:
        import torch.distributed._symmetric_memory as symm_mem
        t = symm_mem.empty(1 << 20, dtype=torch.float32, device=f"cuda:{torch.cuda.current_device()}")
        hdl = symm_mem.rendezvous(t, dist.group.WORLD.group_name)
        out["buffer_ptrs"] = [hex(p) for p in list(hdl.buffer_ptrs)[:world]]
        out["multicast_ptr"] = hex(int(hdl.multicast_ptr))
        out["has_multicast"] = int(hdl.multicast_ptr) != 0
        t.fill_(float(rank + 1))
        hdl.barrier()
        try:
            torch.ops.symm_mem.multimem_all_reduce_(t, "sum", dist.group.WORLD.group_name)
            torch.cuda.synchronize()
            out["multimem_all_reduce"] = t[:2].tolist()
            out["multimem_all_reduce_ok"] = abs(t[0].item() - world * (world + 1) / 2) < 1e-6
        except Exception as e:  # noqa: BLE001
            out["multimem_all_reduce_error"] = f"{type(e).__name__}: {e}"[:300]
    except Exception as e:  # noqa: BLE001
        out["error"] = f"{type(e).__name__}: {e}"[:400]
        out["trace"] = traceback.format_exc()[-600:]
    print("NVLS_PROBE", out, flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
