#!/usr/bin/env python
"""Does this box offer NVLink-SHARP multicast (NVLS) to user kernels?  Launch with torchrun, one rank per GPU.

Reports (rank 0, JSON): the CUDA driver's multicast attribute, whether torch's symmetric-memory rendezvous returns a
multicast address, and the peer pointers it returns.  NCCL's own view is visible with NCCL_DEBUG=INFO (grep NVLS).
"""
import json
import os
import sys

import torch
import torch.distributed as dist


def main() -> int:
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    out = {"world": world}
    try:
        from cuda.bindings import driver as cu
        cu.cuInit(0)
        err, dev = cu.cuDeviceGet(local)
        err, v = cu.cuDeviceGetAttribute(cu.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev)
        out["driver_multicast_supported"] = int(v)
    except Exception as e:  # noqa: BLE001
        out["driver_multicast_supported"] = f"query failed: {e!r}"
    try:
        import torch.distributed._symmetric_memory as symm
        t = symm.empty(1 << 20, dtype=torch.float32, device=torch.device("cuda", local))
        h = symm.rendezvous(t, dist.group.WORLD)
        out["symm_backend"] = symm.get_backend(torch.device("cuda", local)) if hasattr(symm, "get_backend") else "?"
        out["multicast_ptr"] = int(getattr(h, "multicast_ptr", 0))
        out["buffer_ptrs"] = [int(p) for p in h.buffer_ptrs]
        out["signal_pad_ptrs"] = [int(p) for p in h.signal_pad_ptrs]
        if out["multicast_ptr"]:
            # functional check with torch's own multimem all-reduce
            t.fill_(float(rank + 1))
            dist.barrier()
            torch.ops.symm_mem.multimem_all_reduce_(t, "sum", dist.group.WORLD.group_name)
            torch.cuda.synchronize()
            out["multimem_all_reduce_value"] = float(t[0].item())
            out["expected"] = world * (world + 1) / 2
    except Exception as e:  # noqa: BLE001
        out["symm_error"] = repr(e)
    x = torch.ones(1 << 22, device="cuda")
    dist.all_reduce(x)
    torch.cuda.synchronize()
    if rank == 0:
        print("NVLS_PROBE " + json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
