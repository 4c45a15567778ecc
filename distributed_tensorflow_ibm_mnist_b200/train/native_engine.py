"""The B200 execution engine: one training step = a fixed sequence of hand-written sm_100a kernels, replayed as a
CUDA graph, with every piece of state (weights, optimizer slots, EMA shadows, global_step, loss averages) resident
on the device.

Replaces what the reference delegates to TensorFlow's executor inside ``mon_sess.run(train_op)`` (main.py:148-151):
forward (mnist_input.py:118-208), loss/accuracy (:210-243), autodiff, optimizer apply + EMA + global_step (:246-270).
Step schedule (``baseline`` preset; ``reference`` adds LRN inside the conv kernels, the local4 GEMMs and a separate
un-pooling kernel).  ``||`` = concurrent branches on forked streams; consecutive kernels are linked by programmatic
dependent launch edges, so each prologue overlaps its predecessor's tail (DESIGN.md section 2 has the measured timeline):

    conv1[tcgen05] -> conv2_fwd[tcgen05] -> fc GEMM[tcgen05+TMA] -> head_forward (loss, dlogits, dH)            forward
    dP2 GEMM (+ max-pool backward in its epilogue)  ||  head_backward  ||  dW3 GEMM                              fc backward
    conv2_dgrad -> conv2_wgrad || conv1_wgrad         ||  optimizer update of the fc/head bucket (one CTA per SM)   conv backward
    optimizer update of the conv bucket + step bookkeeping                                                       close
  (with > 1 rank both updates are the fused peer-memory exchange: reduce-scatter + optimizer + all-gather in one kernel)
"""
from __future__ import annotations

import os

import torch

from ..models.params import FlatParams
from ..models.presets import ModelSpec
from ..ops import layouts, native


class NativeEngine:
    def __init__(self, spec: ModelSpec, batch_size: int, *, device="cuda", optimizer="adam", base_lr=1e-3, decay_rate=0.1,
                 decay_steps=0, momentum=0.9, seed=0, comm=None, use_graph=True, keep_logits=False):
        if batch_size % 8 != 0:
            raise ValueError("the native engine needs a batch size that is a multiple of 8 (TMA row alignment)")
        if not torch.cuda.is_available():
            raise RuntimeError("NativeEngine requires a CUDA device (sm_100a); use TorchEngine on CPU")
        from ..ops import _lib

        _lib.load()  # fail loudly if the kernel library is missing
        self.spec, self.B = spec, batch_size
        self.device = torch.device(device)
        self.comm = comm
        self.world = comm.world_size if comm is not None else 1
        self.seed = seed
        self.use_graph = use_graph
        self.optimizer = optimizer
        alloc = comm.flat_allocator() if comm is not None else None
        self.flat = FlatParams(spec, device=self.device, allocator=alloc)
        self.cfg = self.flat.update_config(optimizer=optimizer, base_lr=base_lr, decay_rate=decay_rate,
                                           decay_steps=decay_steps, momentum=momentum)
        self.state = native.new_state(self.device)
        native.state_init(self.state, 0, self.cfg.beta1, self.cfg.beta2)
        self.eval_state = native.new_state(self.device)
        native.state_init(self.eval_state)
        self._alloc_activations(keep_logits)
        # the update kernels keep the conv2 operand images in step with the weights (dtm_update_config.pack_*)
        self.cfg.pack_seg = self.flat.index["conv2/weights"]
        self.cfg.pack_fwd = self.w2_fwd.data_ptr()
        self.cfg.pack_dgrad = self.w2_dgrad.data_ptr()
        # single GPU: the whole optimizer state (72 MB for the baseline preset) lives in the persisting part of the 126 MB L2
        # (measured, same box: 159.8 -> 157.6 us per step with the separate weight-gradient kernels, 162.4 -> 152.6 with the
        # fused one; DTM_L2_PERSIST=0 switches it off)
        self.l2_window = None
        if self.flat.blob is not None and os.environ.get("DTM_L2_PERSIST", "1") != "0":
            try:
                self.l2_window = native.set_update_l2_window(self.flat.blob)
                if self.l2_window["carveout_bytes"] < self.l2_window["bytes"]:
                    # the state does not fit the persisting carve-out (wide preset: 568 MB): a partial window only makes
                    # the update kernels thrash the lines they pinned (measured: 497 -> 715 us per step) - switch it off
                    native.set_update_l2_window(None)
                    self.l2_window = None
            except _lib.NativeError:
                self.l2_window = None
        self._graphs: dict[tuple, torch.cuda.CUDAGraph] = {}
        self._side = torch.cuda.Stream(device=self.device)
        self._comm_stream = torch.cuda.Stream(device=self.device)
        self._head_stream = torch.cuda.Stream(device=self.device)
        self._schedule_policy(comm)
        self._fork_ev = [torch.cuda.Event() for _ in range(10)]
        # gradient buckets in flat-buffer order: the fc/head variables (93 % of the bytes, ready right after the first
        # backward GEMMs) and the conv variables (ready last).  With >1 rank bucket 0 is reduced + updated on its own
        # stream while the conv backward still runs.
        split = self.flat.offsets["local3/weights"]
        self.buckets = [(split, self.flat.total), (0, split)]
        # Multi-GPU: the conv bucket closes the step, so its exchange is pure latency.  Default: replicated one-shot update
        # (every rank sums all ranks' conv gradients itself: one barrier, no all-gather) with the bucket's gradients
        # double-buffered by step parity; DTM_CONV_EXCHANGE=fused keeps the reduce-scatter / all-gather kernel for it.
        self._grad_base = None          # summaries: gradients of an extra pass go to a private buffer
        self._parity = 0
        self.grad_alt = None
        self._replicated_conv = self.sharded and os.environ.get("DTM_CONV_EXCHANGE", "replicated") != "fused"
        if self._replicated_conv:
            self._grad_alt_buf = comm.alloc_symmetric(split * 4)
            self.grad_alt = self._grad_alt_buf.view(torch.float32, split)
        self.launches_per_step = 0
        self.trace = None   # set to a list to collect (label, cuda event) marks on the main stream (tools/phase_times.py)

    def _schedule_policy(self, comm) -> None:
        """Measured scheduling choices (tools/ablate.py, tools/timeline.py; numbers in DESIGN.md section 2).  They depend on
        what bounds the step: a chain of 10-30 us kernels for the small presets, dW3 -> optimizer update for ``wide``."""
        spec, env = self.spec, os.environ
        small = self.flat.total <= (8 << 20)
        self._main_priority = -1 if env.get("DTM_PRIORITY", "0") != "0" else 0     # measured harmful with PDL: off
        # dH inside head_forward pays off while the head's weight rows fit one shared-memory staging chunk (1024 hidden
        # units); beyond that (wide) the forward kernel would stage every chunk twice
        self._fuse_dh = (spec.fc2 or spec.fc1) <= 1024
        # programmatic dependent launch ("early weights" mode): hiding the launch gaps of the small presets is worth 10 %;
        # with the wide preset early-resident CTAs of the main chain take SMs away from the kernels that bound the step
        # (536 -> 562 us), so it is off there
        self._pdl = 2 if small else 0
        # grid cap of the early optimizer update: one CTA per SM.  conv2_dgrad is held to 128 registers so that exactly one
        # update CTA fits beside it on every SM - the two kernels run side by side instead of taking turns (179.6 -> 169.6
        # us).  The LRN preset's backward kernels carry heavier epilogues and its fc bucket is small: two update CTAs per SM
        # finish before conv2_dgrad needs the SMs (same box, us per step: uncapped 230.5, 148 CTAs 231, 222: 217.2,
        # 296: 217.0, 444: 227.5, 592: 225.5).  No cap for wide (the update is the long pole).
        sm_count = torch.cuda.get_device_properties(self.device).multi_processor_count
        default_cap = (2 * sm_count if spec.lrn else sm_count) if small else 0
        self._early_update_ctas = int(env.get("DTM_EARLY_UPDATE_CTAS", str(default_cap)))
        if comm is not None and "DTM_OVERLAP_CTAS" not in env:
            if not small:
                # wide: the fc bucket's exchange runs ~200 us and is bound by the NVSwitch, not by the CTA count (8-GPU
                # sweep: 32 CTAs reach the time of 148) - leave most SMs entirely to the conv backward
                comm.overlap_ctas = 64
            else:
                # the overlapped exchange shares its SMs with the conv backward: give it only as many CTAs as the rank's
                # shard needs (about six 256-thread visits per CTA).  8 GPUs, same box: 148 CTAs 167.7 us per step,
                # 64 CTAs 153.4, 32 CTAs 163.0; two ranks (four times the shard) want every SM
                shard4 = (self.flat.total - self.flat.offsets["local3/weights"]) // self.world // 4
                comm.overlap_ctas = max(32, min(sm_count, -(-shard4 // (256 * 6))))
        # one tcgen05 kernel for both convolution weight gradients (conv1's operands are built by the warps that idle in
        # conv2_wgrad): single-channel uint8 input only; DTM_WGRAD_FUSED=0 restores the two separate kernels.  Pays off
        # together with the persisting-L2 optimizer state (profiles/ab_wgrad_update_r2.md): 152.6 vs 157.6 us per step
        self._fused_wgrad = env.get("DTM_WGRAD_FUSED", "1") != "0"
        # experiment switches (tools/ablate.py): kernels to leave out of the step (timing only - results are then wrong) and
        # a flag that puts every kernel on the main stream
        self._skip = set(filter(None, env.get("DTM_SKIP", "").split(",")))
        self._serial = env.get("DTM_SERIAL", "0") == "1"

    # ------------------------------------------------------------------------------------------------
    def _alloc_activations(self, keep_logits: bool) -> None:
        B, s, dev = self.B, self.spec, self.device
        z = lambda *shape, dtype: torch.zeros(*shape, dtype=dtype, device=dev)
        # one buffer per slot: pixels, then labels - a packed pinned host batch arrives with a single H2D copy
        nb = B * s.image_pixels
        pad = (-nb) % 16
        self.xy = [z(nb + pad + 4 * B, dtype=torch.uint8) for _ in range(2)]
        self.x = [b[:nb].view(B, s.image_pixels) for b in self.xy]
        self.y = [b[nb + pad:].view(torch.int32) for b in self.xy]
        self.packed_bytes, self.packed_label_offset = nb + pad + 4 * B, nb + pad
        self.xf = None  # float input staging for inference with pre-normalised images
        self.p1 = layouts.alloc_planes(B, 32, dev)
        self.arg1 = z(B, 196, 32, dtype=torch.uint8)
        self.pool_raw = z(B, 196, 32, dtype=torch.bfloat16) if s.lrn else None
        self.p2 = z(B, 3136, dtype=torch.bfloat16)
        self.arg2 = z(B, 49, 64, dtype=torch.uint8)
        self.act2 = z(B, 196, 64, dtype=torch.bfloat16) if s.lrn else None
        self.h1 = z(B, s.fc1, dtype=torch.bfloat16)
        self.h2 = z(B, s.fc2, dtype=torch.bfloat16) if s.fc2 else None
        self.logits = z(B, 10, dtype=torch.float32)
        self.probs = z(B, 10, dtype=torch.float32)
        self.dlogits = z(B, 10, dtype=torch.float32)
        self.dh_last = z(B, s.head_in, dtype=torch.bfloat16)
        self.dh1 = z(B, s.fc1, dtype=torch.bfloat16) if s.fc2 else None
        self.dp2 = z(B, 3136, dtype=torch.bfloat16)
        self.dz2 = layouts.alloc_planes(B, 64, dev)
        self.dpool1 = z(B, 196, 32, dtype=torch.bfloat16)
        self.w2_fwd = z(25 * 32 * 64, dtype=torch.bfloat16)      # conv2 filter, packed as tcgen05 operand images
        self.w2_dgrad = z(25 * 32 * 64, dtype=torch.bfloat16)

    # ---- parameters / state ---------------------------------------------------------------------------
    def load_params(self, tensors, *, ema=None, slots=None, global_step=0, loss_avg=None) -> None:
        self.flat.load_dict({k: torch.as_tensor(v) for k, v in tensors.items()}, reset_ema=ema is None)
        if ema:
            self.flat.load_ema_dict(ema)
        if slots:
            for k, d in enumerate(slots):
                if d:
                    self.flat.load_slot_dict(k, d)
        native.conv2_pack_weights(self.flat.weight_bf16("conv2/weights"), self.w2_fwd, self.w2_dgrad)
        native.state_init(self.state, global_step, self.cfg.beta1, self.cfg.beta2)
        if loss_avg:
            st = native.read_state(self.state)
            for i, v in enumerate(loss_avg):
                st.loss_avg[i] = v
            self.state.copy_(torch.frombuffer(bytearray(bytes(st)), dtype=torch.uint8).to(self.device))
        torch.cuda.synchronize(self.device)

    def broadcast_params(self, root: int = 0) -> None:
        """Make rank ``root``'s weights and training-state scalars authoritative on every rank (peer-memory broadcast) and
        refresh everything derived from them (bf16 shadow, conv2 operand images, EMA shadows of a fresh run)."""
        if self.comm is None or self.world <= 1:
            return
        self.comm.broadcast_params(self.flat, self.state, root)
        native.conv2_pack_weights(self.flat.weight_bf16("conv2/weights"), self.w2_fwd, self.w2_dgrad)
        if int(self.read_state().global_step) == 0 and self.flat.ema is not None:
            self.flat.ema.copy_(self.flat.w)     # TF initialises the shadows to the (now common) initial values
        elif self.flat.ema is not None:
            # a restored run: optimizer slots and EMA shadows must be the root's as well - only the root may have seen the
            # checkpoint (train_dir on a local disk), and a rank with fresh slots would silently corrupt its shard
            for buf in (self.flat.slot0, self.flat.slot1, self.flat.ema):
                if buf is not None:
                    self.comm.plane.broadcast_tensor(buf, root)
        torch.cuda.synchronize(self.device)

    @property
    def sharded(self) -> bool:
        """True when optimizer slots / EMA shadows are only valid on their owning rank (fused multi-GPU update)."""
        return self.comm is not None and self.world > 1

    def params(self) -> dict[str, torch.Tensor]:
        # sharded engines keep the fp32 master of the big fc matrices on the owning rank only (collective gather)
        return self._full(self.flat.w) if self.sharded else {k: v.cpu() for k, v in self.flat.to_dict().items()}

    def _full(self, buf: torch.Tensor) -> dict[str, torch.Tensor]:
        if self.sharded:   # collective: every rank must call
            buf = self.comm.gather_sharded(buf, self.buckets)
        return {v.name: buf[self.flat.offsets[v.name]:self.flat.offsets[v.name] + v.size].view(v.shape).cpu() for v in self.spec.variables}

    def ema_params(self) -> dict[str, torch.Tensor]:
        return self._full(self.flat.ema)

    def slots(self) -> list[dict[str, torch.Tensor]]:
        n = {"sgd": 0, "momentum": 1, "adam": 2}[self.optimizer.lower()]
        return [self._full(self.flat.slot0 if i == 0 else self.flat.slot1) for i in range(n)]

    def _g(self, name: str) -> torch.Tensor:
        """Where the backward pass accumulates the gradient of ``name``: the flat gradient buffer, the other half of the
        double-buffered conv bucket on odd steps, or the private buffer of a summary pass."""
        f = self.flat
        v, o = self.spec.var(name), f.offsets[name]
        if self._grad_base is not None:
            buf = self._grad_base
        elif self._parity and self.grad_alt is not None and o < self.buckets[1][1]:
            buf = self.grad_alt
        else:
            buf = f.grad
        return buf[o:o + v.size].view(v.shape)

    def gradients(self) -> dict[str, torch.Tensor]:
        return {k: v.cpu() for k, v in self.flat.grad_dict().items()}

    def read_state(self) -> native.TrainState:
        return native.read_state(self.state)

    @property
    def state_ptr_step(self) -> int:
        return self.state.data_ptr()  # global_step is the first field of dtm_train_state

    def _on(self, name: str) -> bool:
        return name not in self._skip

    def _stream(self, st):
        """Side stream, or the current one when the serial experiment switch is set."""
        return torch.cuda.current_stream() if self._serial else st

    def _mark(self, label: str) -> None:
        if self.trace is not None:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record(torch.cuda.current_stream())
            self.trace.append((label, ev))

    # ---- the kernel sequence ----------------------------------------------------------------------------
    def _forward(self, x, labels, *, training: bool, state, count=None) -> None:
        # early-weights PDL mode: in this schedule no kernel that stages weights before its dependency wait (conv2_fwd,
        # conv2_dgrad, head_forward) directly follows a kernel that writes weights
        with native.pdl_mode(self._pdl):
            self._forward_impl(x, labels, training=training, state=state, count=count)

    def _forward_impl(self, x, labels, *, training: bool, state, count=None) -> None:
        s, f, B = self.spec, self.flat, self.B
        n = 0
        self._mark("start")
        if self._on("conv1_fwd"):
            native.conv1_fwd(x, f.weight("conv1/weights"), f.weight("conv1/biases"), self.p1, self.arg1, self.pool_raw,
                             channels=s.channels, lrn=s.lrn); n += 1
        self._mark("conv1_fwd")
        if self._on("conv2_fwd"):
            native.conv2_fwd(self.p1, f.weight_bf16("conv2/weights"), f.weight("conv2/biases"), self.p2, self.arg2, self.act2,
                             batch=B, lrn=s.lrn, packed=self.w2_fwd); n += 1
        self._mark("conv2_fwd")
        drop_here = training and s.keep_prob < 1.0 and not s.fc2
        if self._on("fc1_fwd"):
            native.gemm_bf16(self.p2, 0, f.weight_bf16("local3/weights"), 1, self.h1, B, s.fc1, 3136,
                             epilogue=native.EPI_BIAS_RELU_BF16, bias=f.weight("local3/biases"),
                             keep_prob=s.keep_prob if drop_here else 1.0, seed=self.seed, step_ptr=self.state_ptr_step); n += 1
        self._mark("fc1_fwd")
        h = self.h1
        if s.fc2:
            native.gemm_bf16(self.h1, 0, f.weight_bf16("local4/weights"), 1, self.h2, B, s.fc2, s.fc1,
                             epilogue=native.EPI_BIAS_RELU_BF16, bias=f.weight("local4/biases")); n += 1
            h = self.h2
        if self._on("head_fwd"):
            # training: the head also emits dH, so the fc gradient GEMMs can follow it directly (head_backward then only
            # produces weight / bias gradients, beside those GEMMs)
            native.head_forward(h, f.weight("softmax_linear/weights"), f.weight("softmax_linear/biases"), labels,
                                grad_scale=1.0 / B, loss_scale=1.0 / B, logits=self.logits, probs=self.probs,
                                dlogits=self.dlogits if training else None, state=state,
                                dh=self.dh_last if (training and labels is not None and self._fuse_dh) else None,
                                inv_keep=1.0 / s.keep_prob if drop_here else 1.0); n += 1
        self._mark("head_fwd")
        if count is not None:
            count[0] += n

    def _backward(self, x, count, with_comm: bool = True) -> None:
        with native.pdl_mode(self._pdl):
            self._backward_impl(x, count, with_comm)

    def _backward_impl(self, x, count, with_comm: bool = True) -> None:
        s, f, B = self.spec, self.flat, self.B
        n = 0
        h = self.h2 if s.fc2 else self.h1
        inv_keep = 1.0 / s.keep_prob if (s.keep_prob < 1.0 and not s.fc2) else 1.0
        main = torch.cuda.current_stream()
        if self._on("head_bwd") and not self._fuse_dh:
            native.head_backward(h, f.weight("softmax_linear/weights"), self.dlogits, self.dh_last,
                                 self._g("softmax_linear/weights"), self._g("softmax_linear/biases"),
                                 self._g("local4/biases" if s.fc2 else "local3/biases"), inv_keep=inv_keep); n += 1
        elif self._on("head_bwd"):
            self._fork_ev[6].record(main)   # the head's own gradients depend on head_forward only; launched further down
        self._mark("head_bwd")
        if s.fc2:
            # dW4 = h1^T . dY2 ; dY1 = (dY2 . W4^T) * [h1 > 0] ; db3 = colsum(dY1)
            native.gemm_bf16(self.h1, 1, self.dh_last, 1, self._g("local4/weights"), s.fc1, s.fc2, B); n += 1
            native.gemm_bf16(self.dh_last, 0, f.weight_bf16("local4/weights"), 0, self.dh1, B, s.fc1, s.fc2,
                             epilogue=native.EPI_MASK_BF16, ref=self.h1); n += 1
            native.colsum_bf16(self.dh1, self._g("local3/biases")); n += 1
            dy1 = self.dh1
        else:
            dy1 = self.dh_last
        lo0, hi0 = self.buckets[0]
        # Three independent branches hang off this point: dP2 (feeds the whole conv backward: critical), dW3 and the head's
        # weight/bias gradients (both only feed the optimizer).  Captured as dP2, head, dW3 - the block scheduler serves concurrently
        # runnable kernels oldest first: with dW3 ahead of it dP2 started 10 us late, with the (light) head kernel last the
        # optimizer update that waits for it started 14 us late (tools/timeline.py).
        self._fork_ev[0].record(main)
        # without LRN the max-pool backward of conv2 rides in the epilogue of this GEMM (one N tile = one pooled pixel)
        fused_unpool = not s.lrn and self._on("fused_unpool")
        if self._on("dP2"):
            if fused_unpool:
                native.gemm_mask_unpool(dy1, f.weight_bf16("local3/weights"), self.p2, self.arg2, self.dp2, self.dz2,
                                        batch=B, hidden=s.fc1); n += 1
            else:
                native.gemm_bf16(dy1, 0, f.weight_bf16("local3/weights"), 0, self.dp2, B, 3136, s.fc1,
                                 epilogue=native.EPI_MASK_BF16, ref=self.p2); n += 1
        if self._on("head_bwd") and self._fuse_dh:
            self._head_stream.wait_event(self._fork_ev[6])
            with torch.cuda.stream(self._stream(self._head_stream)):
                native.head_backward(h, f.weight("softmax_linear/weights"), self.dlogits, self.dh_last,
                                     self._g("softmax_linear/weights"), self._g("softmax_linear/biases"),
                                     self._g("local4/biases" if s.fc2 else "local3/biases"), inv_keep=inv_keep,
                                     dh_ready=True); n += 1
                self._fork_ev[7].record(self._head_stream)
        self._side.wait_event(self._fork_ev[0])
        with torch.cuda.stream(self._stream(self._side)):
            if self._on("dW3"):
                native.gemm_bf16(self.p2, 1, dy1, 1, self._g("local3/weights"), 3136, s.fc1, B); n += 1
            self._fork_ev[1].record(self._side)
        self._mark("dP2_gemm")
        if fused_unpool:
            # conv2 bias gradient = column sums of dP2 over images and pooled pixels: beside the conv backward
            self._fork_ev[8].record(main)
            self._head_stream.wait_event(self._fork_ev[8])
            with torch.cuda.stream(self._stream(self._head_stream)):
                native.colsum_bf16(self.dp2.view(B * 49, 64), self._g("conv2/biases")); n += 1
                self._fork_ev[9].record(self._head_stream)
        # bucket 0 (fc + head variables, 98 % of the elements): every gradient is complete once the dW GEMM is done and the
        # last reader of the fc weights (the dP2 GEMM above) has finished -> update it now, on its own stream, beside the
        # conv backward (tensor-core / latency bound kernels that leave the memory system idle).
        early_update = with_comm and self._on("early_update")
        if early_update:
            self._fork_ev[4].record(main)
            self._comm_stream.wait_event(self._fork_ev[4])
            self._comm_stream.wait_event(self._fork_ev[1])
            if self._on("head_bwd") and self._fuse_dh:
                self._comm_stream.wait_event(self._fork_ev[7])
            with torch.cuda.stream(self._stream(self._comm_stream)):
                if self.sharded:
                    self.comm.fused_allreduce_update(self.cfg, self.state, f, lo0, hi0, finish=False, overlapped=True)
                elif self._on("update0"):
                    # capped grid: the kernel trickles along in the background instead of occupying every SM slot and
                    # stalling the conv kernels queued behind it
                    native.fused_update_range(self.cfg, self.state, f.w, f.grad, f.slot0, f.slot1, f.ema, f.w_bf16, lo0, hi0,
                                              finish=False, max_ctas=self._early_update_ctas)
                n += 1
                self._fork_ev[5].record(self._comm_stream)
        if self._on("unpool2") and not fused_unpool:
            native.unpool2(self.dp2, self.arg2, self.act2, self.dz2, self._g("conv2/biases"), batch=B, lrn=s.lrn); n += 1
        self._mark("unpool2")
        # dgrad first: it and wgrad both need > 113 KB of shared memory and cannot share an SM, so launching them together
        # only serialises them.  conv1_wgrad (26 KB, CUDA cores) does co-reside with the tensor-core wgrad.
        if self._on("conv2_dgrad"):
            native.conv2_dgrad(self.dz2, f.weight_bf16("conv2/weights"), self.p1, self.pool_raw, self.dpool1, batch=B, lrn=s.lrn,
                               packed=self.w2_dgrad); n += 1
        self._mark("conv2_dgrad")
        fused_wgrad = self._fused_wgrad and x.dtype == torch.uint8 and self._on("conv2_wgrad") and self._on("conv1_wgrad")
        if fused_wgrad:
            native.conv_wgrad_fused(self.p1, self.dz2, x, self.dpool1, self.arg1, self._g("conv2/weights"),
                                    self._g("conv1/weights"), self._g("conv1/biases"), batch=B, channels=s.channels); n += 1
        else:
            # conv2_wgrad is the longer of the two weight-gradient kernels: it stays on the main stream (programmatic edge
            # from dgrad, no cross-stream hand-off on the critical path); conv1_wgrad (26 KB of smem, CUDA cores) co-resides
            self._fork_ev[2].record(main)
            self._side.wait_event(self._fork_ev[2])
            with torch.cuda.stream(self._stream(self._side)):
                if self._on("conv1_wgrad"):
                    native.conv1_wgrad(x, self.dpool1, self.arg1, self._g("conv1/weights"), self._g("conv1/biases"),
                                       channels=s.channels); n += 1
                self._fork_ev[3].record(self._side)
            if self._on("conv2_wgrad"):
                native.conv2_wgrad(self.p1, self.dz2, self._g("conv2/weights"), batch=B); n += 1
        self._mark("conv1_wgrad")
        main.wait_event(self._fork_ev[1])
        if not fused_wgrad:
            main.wait_event(self._fork_ev[3])
        if self._on("head_bwd") and self._fuse_dh:
            main.wait_event(self._fork_ev[7])
        if fused_unpool:
            main.wait_event(self._fork_ev[9])
        self._mark("join_wgrad")
        if early_update:
            main.wait_event(self._fork_ev[5])
        self._mark("join_update0")
        self._early_update_done = early_update
        count[0] += n

    def _update(self, count) -> None:
        """Close the step: the conv bucket (bucket 0 was launched from inside the backward pass)."""
        f = self.flat
        lo, hi = self.buckets[1] if getattr(self, "_early_update_done", False) else (0, f.total)
        if not self._on("update1"):
            return
        if self.sharded and self._replicated_conv and getattr(self, "_early_update_done", False):
            now, prev = (self._grad_alt_buf, f.grad) if self._parity else (self.comm.grad_buf, self.grad_alt)
            self.comm.replicated_update(self.cfg, self.state, f, lo, hi, now, prev, finish=True)
        elif self.sharded:
            self.comm.fused_allreduce_update(self.cfg, self.state, f, lo, hi, finish=True)
        else:
            native.fused_update_range(self.cfg, self.state, f.w, f.grad, f.slot0, f.slot1, f.ema, f.w_bf16, lo, hi, finish=True)
        self._mark("update_tail")
        count[0] += 1

    def _step_body(self, slot: int) -> int:
        count = [0]
        self._forward(self.x[slot], self.y[slot], training=True, state=self.state, count=count)
        self._backward(self.x[slot], count)
        self._update(count)
        return count[0]

    # ---- public API -------------------------------------------------------------------------------------
    def stage_batch(self, slot: int, images_u8: torch.Tensor, labels: torch.Tensor, stream=None) -> None:
        """Asynchronous copy of a host (ideally pinned) or device batch into device slot ``slot``."""
        with torch.cuda.stream(stream if stream is not None else torch.cuda.current_stream()):
            self.x[slot].copy_(images_u8.view(self.B, -1), non_blocking=True)
            self.y[slot].copy_(labels.view(self.B).to(torch.int32) if labels.dtype != torch.int32 else labels.view(self.B),
                               non_blocking=True)

    def stage_packed(self, slot: int, packed_u8: torch.Tensor, stream=None) -> None:
        """One asynchronous copy of a packed batch (``packed_bytes`` uint8: pixels, padding to 16 bytes, int32 labels)."""
        with torch.cuda.stream(stream if stream is not None else torch.cuda.current_stream()):
            self.xy[slot].copy_(packed_u8, non_blocking=True)

    def train_step(self, slot: int = 0) -> None:
        """Run one optimisation step on the batch staged in ``slot`` (asynchronous; no host sync)."""
        self._last_slot = slot
        key = (slot, self._parity)
        if not self.use_graph:
            self.launches_per_step = self._step_body(slot)
            if self._replicated_conv:
                self._parity ^= 1
            return
        g = self._graphs.get(key)
        if g is None:
            # warm-up outside capture (sets kernel attributes), then undo its effect on the state by re-capturing
            snap = self._snapshot()
            self.launches_per_step = self._step_body(slot)
            torch.cuda.synchronize(self.device)
            self._restore(snap)
            # Towers that are threads of one process go through the capture in lock step: entering a capture empties torch's
            # allocator cache on EVERY device of the process (cudaFree synchronises those devices), and a sibling that has
            # already replayed its graph would sit in an exchange kernel waiting for this tower - which then could never
            # finish its capture.  Both barriers see idle devices.
            plane = getattr(self.comm, "plane", None)
            lockstep = plane is not None and plane.same_process
            if lockstep:
                plane.barrier()
            g = torch.cuda.CUDAGraph()
            # the critical path (main stream) gets the higher priority: when its kernels and the side-stream kernels (dW3,
            # conv1_wgrad, the early optimizer update) compete for SMs, the block scheduler serves the main chain first
            cap = torch.cuda.Stream(device=self.device, priority=self._main_priority)
            cap.wait_stream(torch.cuda.current_stream())
            with torch.cuda.graph(g, stream=cap, capture_error_mode="thread_local"):   # sibling tower threads keep working
                self._step_body(slot)
            torch.cuda.current_stream().wait_stream(cap)
            self._restore(snap)
            if lockstep:
                plane.barrier()
            self._graphs[key] = g
        g.replay()
        if self._replicated_conv:
            self._parity ^= 1

    def _snapshot(self):
        f = self.flat
        bufs = [f.w, f.grad, f.slot0, f.slot1, f.ema, f.w_bf16, self.state, self.w2_fwd, self.w2_dgrad, self.grad_alt]
        return [b.clone() if b is not None else None for b in bufs]

    def _restore(self, snap) -> None:
        f = self.flat
        bufs = [f.w, f.grad, f.slot0, f.slot1, f.ema, f.w_bf16, self.state, self.w2_fwd, self.w2_dgrad, self.grad_alt]
        for b, s in zip(bufs, snap):
            if b is not None:
                b.copy_(s)
        torch.cuda.synchronize(self.device)

    def forward_backward(self, slot: int = 0) -> None:
        """Forward + backward only (gradients left in the flat buffer); used by tests and the summaries."""
        count = [0]
        self._forward(self.x[slot], self.y[slot], training=True, state=self.eval_state, count=count)
        self._backward(self.x[slot], count, with_comm=False)

    def weight_loss(self) -> float:
        """Sum of the ``wd * l2_loss(w)`` terms of the 'losses' collection (mnist_input.py:113) for the current weights.  A
        sharded engine reads the bf16 shadow of the fc matrices (their fp32 masters live on the owning rank only)."""
        total = 0.0
        for v in self.spec.variables:
            if v.wd:
                w = self.flat.weight_bf16(v.name).float() if (self.sharded and v.tensor_core_only) else self.flat.weight(v.name)
                total += 0.5 * float(v.wd) * float((w * w).sum())
        return total

    def evaluate(self, slot: int = 0) -> tuple[float, float]:
        """(total loss, accuracy) of the staged batch without touching weights - the reference's "test" summaries, which
        fetch the same ``total_loss`` tensor as the train summaries (main.py:103-106): cross-entropy + weight-decay terms."""
        native.state_init(self.eval_state)
        self._forward(self.x[slot], self.y[slot], training=False, state=self.eval_state)
        st = native.read_state(self.eval_state)
        return st.loss_sum + self.weight_loss(), st.correct

    def predict(self, images: torch.Tensor) -> torch.Tensor:
        """Class probabilities for ``images`` ([n <= B, pixels], uint8 raw or float32 already normalised)."""
        n = images.shape[0]
        if n > self.B:
            raise ValueError("predict: more images than the engine batch size")
        if images.dtype == torch.uint8:
            buf = self.x[0]
        else:
            if self.xf is None:
                self.xf = torch.zeros(self.B, self.spec.image_pixels, dtype=torch.float32, device=self.device)
            buf = self.xf
        buf[:n].copy_(images.view(n, -1), non_blocking=True)
        self._forward(buf, None, training=False, state=None)
        return self.probs[:n].clone()

    def activations(self) -> dict[str, torch.Tensor]:
        """Layer outputs of the last forward pass (device tensors).  The conv layers are fused with their pooling
        (and LRN) stage, so "conv1"/"conv2" are the pooled activations rather than the pre-pool ReLU outputs."""
        out = {"conv1": layouts.planes_to_nhwc(self.p1, 32), "conv2": self.p2.view(self.B, 7, 7, 64), "local3": self.h1}
        if self.h2 is not None:
            out["local4"] = self.h2
        out["softmax_linear"] = self.logits
        return out

    def summary_tensors(self) -> dict:
        """Tensors for the chief's summaries (main.py:95-109).  Gradients are recomputed on the last batch with an extra
        forward/backward pass: the fused update kernel clears accumulated gradients as it consumes them.

        Multi-GPU: only the chief calls this, while its peers have already launched their next step - and a slower peer may
        still be summing the chief's gradients of the step that just ended.  The extra pass therefore accumulates into a
        private gradient buffer, never into the buffers the exchange kernels read."""
        slot = getattr(self, "_last_slot", 0)
        torch.cuda.synchronize(self.device)
        if self.sharded:
            # never the exchanged buffers: a slower peer may still be summing this rank's gradients of the last step
            if getattr(self, "_summary_grad", None) is None:
                self._summary_grad = torch.zeros_like(self.flat.grad)
            self._summary_grad.zero_()
            self._grad_base = self._summary_grad
            try:
                self.forward_backward(slot)
            finally:
                self._grad_base = None
            base = self._summary_grad
            grads = {v.name: base[self.flat.offsets[v.name]:self.flat.offsets[v.name] + v.size].view(v.shape).clone() for v in self.spec.variables}
        else:
            self.flat.grad.zero_()
            self.forward_backward(slot)
            grads = {v.name: self.flat.gradient(v.name).clone() for v in self.spec.variables}
            self.flat.grad.zero_()
        # the bf16 shadow is current on every rank even where the fp32 master lives elsewhere (sharded fc matrices)
        params = {v.name: (self.flat.weight_bf16(v.name).float() if self.sharded and v.tensor_core_only else self.flat.weight(v.name))
                  for v in self.spec.variables}
        return {"params": params, "activations": self.activations(), "gradients": grads}

    def scalar_state(self) -> dict:
        st = self.read_state()
        return {"global_step": int(st.global_step), "beta1_power": float(st.beta1_power), "beta2_power": float(st.beta2_power),
                "loss_avg": [float(v) for v in st.loss_avg]}

    def metrics(self) -> dict:
        st = self.read_state()
        return {"global_step": st.global_step, "cross_entropy": st.last_xent, "total_loss": st.last_total,
                "accuracy": st.last_acc, "lr": st.last_lr, "nan": bool(st.nan_flag)}
