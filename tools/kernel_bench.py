#!/usr/bin/env python
"""Stand-alone device time of the step's kernels (CUDA events around a loop of back-to-back launches, L2-warm), with the
operands a real forward/backward pass left in the engine's buffers.  For A/B work on one kernel - the step time that counts
is bench.py's.   python tools/kernel_bench.py [--preset baseline] [--batch 512] [--only conv_wgrad_fused,...]"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--preset", default="baseline")
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--iters", type=int, default=50)
    ap.add_argument("--only", default="")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    from distributed_tensorflow_ibm_mnist_b200.models import oracle, presets
    from distributed_tensorflow_ibm_mnist_b200.ops import native
    from distributed_tensorflow_ibm_mnist_b200.train.native_engine import NativeEngine

    spec = presets.make_spec(args.preset)
    B = args.batch
    eng = NativeEngine(spec, B, optimizer="adam", use_graph=False)
    eng.load_params(oracle.init_params(spec))
    x = torch.randint(0, 256, (B, spec.image_pixels), dtype=torch.uint8).cuda()
    y = torch.randint(0, 10, (B,), dtype=torch.int32).cuda()
    eng.stage_batch(0, x, y)
    eng.forward_backward(0)
    torch.cuda.synchronize()
    f, s = eng.flat, spec
    xs = eng.x[0]
    lo0, hi0 = eng.buckets[0]
    sm = torch.cuda.get_device_properties(0).multi_processor_count

    def env(**kw):
        class _E:
            def __enter__(self_):
                self_.old = {k: os.environ.get(k) for k in kw}
                os.environ.update({k: str(v) for k, v in kw.items()})
            def __exit__(self_, *a):
                for k, v in self_.old.items():
                    if v is None:
                        os.environ.pop(k, None)
                    else:
                        os.environ[k] = v
        return _E()

    cases = {
        "conv1_fwd": lambda: native.conv1_fwd(xs, f.weight("conv1/weights"), f.weight("conv1/biases"), eng.p1, eng.arg1, eng.pool_raw,
                                              channels=s.channels, lrn=s.lrn),
        "conv2_fwd": lambda: native.conv2_fwd(eng.p1, f.weight_bf16("conv2/weights"), f.weight("conv2/biases"), eng.p2, eng.arg2, eng.act2,
                                              batch=B, lrn=s.lrn, packed=eng.w2_fwd),
        "fc1_fwd": lambda: native.gemm_bf16(eng.p2, 0, f.weight_bf16("local3/weights"), 1, eng.h1, B, s.fc1, 3136,
                                            epilogue=native.EPI_BIAS_RELU_BF16, bias=f.weight("local3/biases"), keep_prob=0.5, seed=1,
                                            step_ptr=eng.state_ptr_step),
        "head_fwd": lambda: native.head_forward(eng.h2 if s.fc2 else eng.h1, f.weight("softmax_linear/weights"), f.weight("softmax_linear/biases"),
                                                eng.y[0], grad_scale=1.0 / B, loss_scale=1.0 / B, logits=eng.logits, probs=eng.probs,
                                                dlogits=eng.dlogits, state=eng.eval_state, dh=eng.dh_last if eng._fuse_dh else None, inv_keep=2.0),
        "head_bwd": lambda: native.head_backward(eng.h2 if s.fc2 else eng.h1, f.weight("softmax_linear/weights"), eng.dlogits, eng.dh_last,
                                                 f.gradient("softmax_linear/weights"), f.gradient("softmax_linear/biases"),
                                                 f.gradient("local4/biases" if s.fc2 else "local3/biases"), inv_keep=2.0, dh_ready=eng._fuse_dh),
        "dW3": lambda: native.gemm_bf16(eng.p2, 1, eng.dh1 if s.fc2 else eng.dh_last, 1, f.gradient("local3/weights"), 3136, s.fc1, B),
        "conv2_dgrad": lambda: native.conv2_dgrad(eng.dz2, f.weight_bf16("conv2/weights"), eng.p1, eng.pool_raw, eng.dpool1, batch=B, lrn=s.lrn,
                                                  packed=eng.w2_dgrad),
        "conv2_wgrad": lambda: native.conv2_wgrad(eng.p1, eng.dz2, f.gradient("conv2/weights"), batch=B),
        "conv1_wgrad": lambda: native.conv1_wgrad(xs, eng.dpool1, eng.arg1, f.gradient("conv1/weights"), f.gradient("conv1/biases"),
                                                  channels=s.channels),
        "update_fc_uncapped": lambda: native.fused_update_range(eng.cfg, eng.eval_state, f.w, f.grad, f.slot0, f.slot1, f.ema, f.w_bf16, lo0, hi0,
                                                                finish=False, max_ctas=0),
        "update_fc_1cta_per_sm": lambda: native.fused_update_range(eng.cfg, eng.eval_state, f.w, f.grad, f.slot0, f.slot1, f.ema, f.w_bf16, lo0,
                                                                   hi0, finish=False, max_ctas=sm),
        "update_conv_tail": lambda: native.fused_update_range(eng.cfg, eng.eval_state, f.w, f.grad, f.slot0, f.slot1, f.ema, f.w_bf16, 0, lo0,
                                                              finish=True),
    }
    if not s.lrn:
        cases["dP2_unpool"] = lambda: native.gemm_mask_unpool(eng.dh_last, f.weight_bf16("local3/weights"), eng.p2, eng.arg2, eng.dp2, eng.dz2,
                                                              batch=B, hidden=s.fc1)
    if True:
        fused = lambda: native.conv_wgrad_fused(eng.p1, eng.dz2, xs, eng.dpool1, eng.arg1, f.gradient("conv2/weights"),
                                                f.gradient("conv1/weights"), f.gradient("conv1/biases"), batch=B, channels=s.channels)
        cases["conv_wgrad_fused"] = fused
        cases["conv_wgrad_fused_no_conv1"] = ("env", {"DTM_FW_NO_CONV1": 1}, fused)
        cases["conv_wgrad_fused_no_c1_mma"] = ("env", {"DTM_FW_NO_CONV1": 3}, fused)
        cases["conv_wgrad_fused_no_c1_build"] = ("env", {"DTM_FW_NO_CONV1": 5}, fused)
        cases["conv_wgrad_fused_c1_sync_only"] = ("env", {"DTM_FW_NO_CONV1": 7}, fused)
    only = set(filter(None, args.only.split(",")))
    rows = {}
    for name, fn in cases.items():
        if only and name not in only:
            continue
        ctx = None
        if isinstance(fn, tuple):
            ctx, fn = env(**fn[1]), fn[2]
            ctx.__enter__()
        for mode in (0, 2):
            with native.pdl_mode(mode):
                for _ in range(5):
                    fn()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(args.iters):
                    fn()
                e1.record()
                torch.cuda.synchronize()
                rows.setdefault(name, {})["pdl%d_us" % mode] = round(e0.elapsed_time(e1) / args.iters * 1e3, 2)
        if ctx is not None:
            ctx.__exit__()
        print(f"{name:28s} {rows[name]}", flush=True)
    if args.out:
        json.dump(rows, open(args.out, "w"), indent=1)
    return 0


if __name__ == "__main__":
    sys.exit(main())
