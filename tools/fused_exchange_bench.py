#!/usr/bin/env python
"""Time the fused gradient exchange (all-reduce + Adam + EMA + all-gather, one kernel) alone, per transport and
in-flight depth.  Launch with torchrun, one rank per GPU.  Device-timed, max over ranks."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tools.multi_gpu_check import timed  # noqa: E402


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    ap.add_argument("--depths", default="1,2,4,8")
    ap.add_argument("--presets", default="baseline,wide")
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from distributed_tensorflow_ibm_mnist_b200.models import presets
    from distributed_tensorflow_ibm_mnist_b200.models.params import FlatParams
    from distributed_tensorflow_ibm_mnist_b200.ops import native
    from distributed_tensorflow_ibm_mnist_b200.parallel.peer_comm import PeerComm

    comm = PeerComm(dist.group.WORLD, dev)
    rows = []
    for preset in args.presets.split(","):
        for use_mc in ([True, False] if comm.nvls else [False]):
            spec = presets.make_spec(preset)
            comm.ctx = None
            comm.use_multicast = use_mc
            flat = FlatParams(spec, device=dev, allocator=comm.flat_allocator())
            cfg = flat.update_config(optimizer="adam", base_lr=1e-3)
            st = native.new_state(dev)
            native.state_init(st)
            for depth in (args.depths.split(",") if use_mc else ["-"]):
                if depth != "-":
                    os.environ["DTM_COMM_DEPTH"] = depth
                ms = timed(lambda: comm.fused_allreduce_update(cfg, st, flat, 0, flat.total, finish=True), dev, 30)
                S = flat.total * 4.0
                f = (world - 1) / world
                row = {"preset": preset, "nvls": use_mc, "depth": depth, "us": round(ms * 1e3, 1),
                       "frac_two_shot_roofline": round(2 * f * S / 900e9 * 1e6 / (ms * 1e3), 3),
                       "frac_nvls_roofline": round((world + 1) / world * S / 900e9 * 1e6 / (ms * 1e3), 3)}
                rows.append(row)
                if rank == 0:
                    print("[fused] " + json.dumps(row), flush=True)
            if use_mc:
                os.environ["DTM_COMM_DEPTH"] = "4"
                for ctas in (32, 64, 96):
                    comm.overlap_ctas = ctas
                    ms = timed(lambda: comm.fused_allreduce_update(cfg, st, flat, 0, flat.total, finish=True, overlapped=True), dev, 30)
                    row = {"preset": preset, "nvls": True, "depth": "4", "ctas": ctas, "us": round(ms * 1e3, 1)}
                    rows.append(row)
                    if rank == 0:
                        print("[fused] " + json.dumps(row), flush=True)
    os.environ.pop("DTM_COMM_DEPTH", None)
    if rank == 0 and args.out:
        json.dump({"world": world, "rows": rows}, open(args.out, "w"), indent=1)
    comm.close()
    dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
