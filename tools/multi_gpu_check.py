#!/usr/bin/env python
"""Multi-GPU correctness + bandwidth harness for the peer-memory communicator (run under torchrun, one rank per GPU).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
        tools/multi_gpu_check.py [--sweep] [--out gpurun_out/multi_gpu_N.json]

Checks: all-reduce kernels value-compared with NCCL and bit-compared with a fixed-order host reduction;
broadcast; the fused all-reduce+update against (NCCL all-reduce -> single-GPU fused update) on the same gradients;
a few real training steps (all ranks must end with identical weights).  --sweep adds the 4 KB - 256 MB bandwidth
sweep against NCCL (BASELINE.json config #5); every timing is CUDA-event based, max over ranks.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def max_over_ranks(ms: float, dev) -> float:
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def timed(fn, dev, iters: int, warmup: int = 3) -> float:
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return max_over_ranks(e0.elapsed_time(e1) / iters, dev)


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--sweep", action="store_true")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from distributed_tensorflow_ibm_mnist_b200.models import oracle, presets
    from distributed_tensorflow_ibm_mnist_b200.models.params import FlatParams
    from distributed_tensorflow_ibm_mnist_b200.ops import native
    from distributed_tensorflow_ibm_mnist_b200.parallel.peer_comm import PeerComm
    from distributed_tensorflow_ibm_mnist_b200.train.native_engine import NativeEngine

    comm = PeerComm(dist.group.WORLD, dev)
    report: dict = {"world": world, "nvls": comm.nvls, "checks": {}, "sweep": []}
    if rank == 0:
        print(f"[info] world={world} nvls_multicast={comm.nvls}", flush=True)
    ok = True

    def check(name: str, cond: bool, info=None):
        nonlocal ok
        flag = torch.tensor([1 if cond else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        good = bool(flag.item())
        ok = ok and good
        if not good:       # which rank, and what it saw
            every = [None] * world
            dist.all_gather_object(every, (bool(cond), repr(info)))
            info = {"per_rank": every}
        report["checks"][name] = {"pass": good, "info": info}
        if rank == 0:
            print(f"[check] {name}: {'PASS' if good else 'FAIL'} {info if info is not None else ''}", flush=True)

    # ---- stand-alone all-reduce: both algorithms, several sizes -----------------------------------------
    for n in (4, 1024, 65536 + 4, 3_276_800):
        g = torch.Generator(device="cpu").manual_seed(1000 * rank + n)
        x = torch.randn(n, generator=g).to(dev)
        want = x.clone()
        dist.all_reduce(want)
        # fixed-order host reduction: rank 0 + rank 1 + ... in fp32
        parts = [torch.empty_like(x) for _ in range(world)]
        dist.all_gather(parts, x)
        fixed = parts[0].clone()
        for p in parts[1:]:
            fixed += p
        for mode in ("one_shot", "two_shot"):
            got = comm.all_reduce(x, mode, nvls=False)
            torch.cuda.synchronize()
            check(f"allreduce_{mode}_{n}_vs_nccl", torch.allclose(got, want, rtol=1e-5, atol=1e-5),
                  float((got - want).abs().max()))
            check(f"allreduce_{mode}_{n}_bit_exact_fixed_order", torch.equal(got, fixed))
            if comm.nvls:
                # in-switch reduction: the summation order is the switch's, so compare numerically and require that
                # every rank holds the same bits
                got = comm.all_reduce(x, mode, nvls=True)
                torch.cuda.synchronize()
                check(f"allreduce_nvls_{mode}_{n}_vs_nccl", torch.allclose(got, want, rtol=1e-5, atol=1e-5),
                      float((got - want).abs().max()))
                g0 = got.clone()
                dist.broadcast(g0, src=0)
                check(f"allreduce_nvls_{mode}_{n}_ranks_agree_bitwise", torch.equal(g0, got))

    # ---- fused all-reduce + update vs NCCL + single-GPU update --------------------------------------------
    for kind in ("sgd", "momentum", "adam"):
        spec = presets.make_spec("reference", channels=1)
        params = oracle.init_params(spec, seed=5)
        flat_a = FlatParams(spec, device=dev, allocator=comm.flat_allocator())
        comm.ctx = None                                  # new symmetric buffers -> new context
        flat_b = FlatParams(spec, device=dev)
        flat_a.load_dict(params)
        flat_b.load_dict(params)
        cfg = flat_a.update_config(optimizer=kind, base_lr=0.05, decay_rate=0.5, decay_steps=2)
        st_a, st_b = native.new_state(dev), native.new_state(dev)
        native.state_init(st_a)
        native.state_init(st_b)
        split = flat_a.offsets["local3/weights"]
        for it in range(3):
            g = torch.Generator(device="cpu").manual_seed(77 * rank + it)
            grad = (torch.randn(flat_a.total, generator=g) * 0.01).to(dev)
            flat_a.grad.copy_(grad)
            red = grad.clone()
            dist.all_reduce(red)
            flat_b.grad.copy_(red)
            comm.fused_allreduce_update(cfg, st_a, flat_a, split, flat_a.total, finish=False)
            comm.fused_allreduce_update(cfg, st_a, flat_a, 0, split, finish=True)
            native.fused_update(cfg, st_b, flat_b.w, flat_b.grad, flat_b.slot0, flat_b.slot1, flat_b.ema, flat_b.w_bf16,
                                grad_scale=1.0 / world)
            torch.cuda.synchronize()
        tol = 3e-4 if kind == "adam" else 2e-6
        buckets = [(split, flat_a.total), (0, split)]
        full_w = comm.gather_sharded(flat_a.w, buckets)       # fp32 masters of the big fc matrices live on their owners
        err = float((full_w - flat_b.w).abs().max() / flat_b.w.abs().max())
        check(f"fused_update_{kind}_weights", err < tol, err)
        check(f"fused_update_{kind}_bf16", torch.equal(flat_a.w_bf16, full_w.to(torch.bfloat16)))
        lo, hi = comm.shard_range(split, flat_a.total, rank)
        e_err = float((flat_a.ema[lo:hi] - flat_b.ema[lo:hi]).abs().max() / flat_b.ema.abs().max())
        check(f"fused_update_{kind}_ema_shard", e_err < tol, e_err)
        sa, sb = native.read_state(st_a), native.read_state(st_b)
        check(f"fused_update_{kind}_state", sa.global_step == sb.global_step == 3 and abs(sa.last_lr - sb.last_lr) < 1e-9 and
              abs(sa.last_wl[4] - sb.last_wl[4]) <= 2e-3 * abs(sb.last_wl[4]) + 1e-9,
              (sa.global_step, sb.global_step, sa.last_wl[4], sb.last_wl[4], sa.last_lr, sb.last_lr))
        zero_ok = float(flat_a.gradient("conv2/weights").abs().sum()) == 0.0 and float(flat_a.gradient("local3/weights").abs().sum()) > 0
        check(f"fused_update_{kind}_grad_clearing", zero_ok)
        w0 = flat_a.w_bf16.clone()
        dist.broadcast(w0, src=0)
        check(f"fused_update_{kind}_ranks_agree_bitwise", torch.equal(w0, flat_a.w_bf16))
        full_ema = comm.gather_sharded(flat_a.ema, buckets)
        check(f"fused_update_{kind}_gather_sharded", float((full_ema - flat_b.ema).abs().max() / flat_b.ema.abs().max()) < tol)

    # ---- broadcast + real training steps through the engine -------------------------------------------------
    spec = presets.make_spec("baseline")
    comm.ctx = None
    eng = NativeEngine(spec, 64, device=dev, optimizer="momentum", base_lr=0.02, seed=11, comm=comm, use_graph=True)
    eng.load_params(oracle.init_params(spec, seed=rank))          # deliberately different per rank
    eng.broadcast_params()
    w0 = eng.flat.w.clone()
    dist.broadcast(w0, src=0)
    check("broadcast_params", torch.equal(w0, eng.flat.w) and torch.equal(eng.flat.w_bf16, eng.flat.w.to(torch.bfloat16)))
    g = torch.Generator(device="cpu").manual_seed(100 + rank)
    x = torch.randint(0, 256, (64, 784), dtype=torch.uint8, generator=g).to(dev)
    y = torch.randint(0, 10, (64,), dtype=torch.int32, generator=g).to(dev)
    eng.stage_batch(0, x, y)
    # cross-check against an un-fused run of the same steps: engine without comm + NCCL all-reduce of the gradients +
    # single-GPU fused update with grad_scale = 1/N.  Different data per rank; compared after the first step (tight: only
    # summation order differs) and after 8 steps (loose: ReLU / dropout decisions amplify rounding differences).
    ref = NativeEngine(spec, 64, device=dev, optimizer="momentum", base_lr=0.02, seed=11, use_graph=False)
    ref.load_params(oracle.init_params(spec, seed=0))
    ref.stage_batch(0, x, y)

    def ref_step():
        count = [0]
        ref._forward(ref.x[0], ref.y[0], training=True, state=ref.state, count=count)
        ref._backward(ref.x[0], count, with_comm=False)
        dist.all_reduce(ref.flat.grad)
        native.fused_update(ref.cfg, ref.state, ref.flat.w, ref.flat.grad, ref.flat.slot0, ref.flat.slot1, ref.flat.ema,
                            ref.flat.w_bf16, grad_scale=1.0 / world)

    def rel_err():
        torch.cuda.synchronize()
        full = comm.gather_sharded(eng.flat.w, eng.buckets)
        return float((full - ref.flat.w).abs().max() / ref.flat.w.abs().max())

    eng.train_step(0)
    ref_step()
    e1 = rel_err()
    check("engine_matches_nccl_path_step1", e1 < 2e-5, e1)
    for _ in range(7):
        eng.train_step(0)
        ref_step()
    e8 = rel_err()
    check("engine_matches_nccl_path_step8", e8 < 5e-2, e8)
    w0 = eng.flat.w_bf16.clone()
    dist.broadcast(w0, src=0)
    m = eng.metrics()
    check("engine_ranks_agree_bitwise", torch.equal(w0, eng.flat.w_bf16), m["global_step"])
    # learning: identical data on every rank (the averaged gradient is then the single-GPU gradient), same path
    g = torch.Generator(device="cpu").manual_seed(4242)
    xs = torch.randint(0, 256, (64, 784), dtype=torch.uint8, generator=g).to(dev)
    ys = torch.randint(0, 10, (64,), dtype=torch.int32, generator=g).to(dev)
    eng.stage_batch(0, xs, ys)
    losses = []
    for _ in range(10):
        eng.train_step(0)
        losses.append(eng.metrics()["cross_entropy"])
    m = eng.metrics()
    check("engine_loss_decreases", losses[-1] < losses[0] and m["global_step"] == 18 and not m["nan"], losses)

    # ---- bandwidth sweep vs NCCL ------------------------------------------------------------------------------
    if args.sweep:
        size = 4 * 1024
        while size <= 256 * 1024 * 1024:
            n = size // 4
            x = torch.ones(n, device=dev)
            buf = comm.scratch(size)
            out = torch.empty(n, device=dev)
            iters = 50 if size <= (1 << 22) else 10
            row = {"bytes": size}
            variants = [("", False)] + ([("nvls_", True)] if comm.nvls else [])
            for prefix, use_mc in variants:
                for mode in ("one_shot", "two_shot"):
                    if mode == "one_shot" and size > (64 << 20):
                        continue
                    ms = timed(lambda: comm.all_reduce_inplace(buf, n, out, mode, use_mc), dev, iters)
                    row[prefix + mode + "_us"] = ms * 1e3
                    row[prefix + mode + "_busbw_GBs"] = 2 * (world - 1) / world * size / (ms * 1e-3) / 1e9
            ms = timed(lambda: dist.all_reduce(x), dev, iters)
            row["nccl_us"] = ms * 1e3
            row["nccl_busbw_GBs"] = 2 * (world - 1) / world * size / (ms * 1e-3) / 1e9
            report["sweep"].append(row)
            if rank == 0:
                print("[sweep] " + json.dumps(row), flush=True)
            size *= 4
        # the fused gradient exchange of the two model sizes: time per call and fraction of the NVLink roofline.
        #   canonical roofline  = 2 (N-1)/N S / 900 GB/s: what a two-shot fp32 all-reduce has to move per direction
        #                         (BASELINE.md section 3);
        #   own-bytes roofline  = the bytes THIS kernel moves in its busier direction / 900 GB/s (bf16-only all-gather of
        #                         the tensor-core-only variables; in-switch reduction when NVLS is on).
        LINK = 900e9
        for preset in ("baseline", "wide"):
            for use_mc in ([False, True] if comm.nvls else [False]):
                spec = presets.make_spec(preset)
                comm.ctx = None
                comm.use_multicast = use_mc
                flat = FlatParams(spec, device=dev, allocator=comm.flat_allocator())
                cfg = flat.update_config(optimizer="adam", base_lr=1e-3)
                st = native.new_state(dev)
                native.state_init(st)
                ms = timed(lambda: comm.fused_allreduce_update(cfg, st, flat, 0, flat.total, finish=True), dev, 20)
                S = flat.total * 4.0
                gather = sum((2.0 if cfg.seg_gather_bf16[i] else 4.0) * (cfg.seg_offset[i + 1] - cfg.seg_offset[i])
                             for i in range(cfg.n_segments))
                f = (world - 1) / world
                if use_mc:
                    out_b, in_b = S + gather / world, S / world + f * gather
                else:
                    out_b = in_b = f * (S + gather)
                own = max(out_b, in_b)
                row = {"fused_allreduce_adam": preset, "nvls": use_mc, "elements": flat.total, "us": ms * 1e3,
                       "canonical_link_bytes_per_dir": 2 * f * S, "canonical_roofline_us": 2 * f * S / LINK * 1e6,
                       "fraction_of_canonical_roofline": 2 * f * S / LINK * 1e6 / (ms * 1e3),
                       "own_link_bytes_busier_dir": own, "own_roofline_us": own / LINK * 1e6,
                       "fraction_of_own_roofline": own / LINK * 1e6 / (ms * 1e3),
                       "effective_link_GBs": own / (ms * 1e-3) / 1e9}
                report["sweep"].append(row)
                if rank == 0:
                    print("[sweep] " + json.dumps(row), flush=True)
        comm.use_multicast = comm.nvls

    report["ok"] = ok
    if rank == 0:
        print("MULTI_GPU_RESULT " + json.dumps({"world": world, "ok": ok, "failed": [k for k, v in report["checks"].items() if not v["pass"]]}))
        if args.out:
            os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
            with open(args.out, "w") as f:
                json.dump(report, f, indent=1)
    comm.close()
    dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
