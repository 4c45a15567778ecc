"""Python wrappers over the sm_100a kernels in libdtm_b200 (one function per C entry point).

Tensors are torch CUDA tensors; every call is enqueued on the current torch CUDA stream, which makes the whole
training step capturable into a CUDA graph.  Layout helpers for the "plane layout" live in :mod:`.layouts`.
"""
from __future__ import annotations

import ctypes
from ctypes import c_float, c_int, c_longlong, c_uint64, c_void_p

import torch

from . import _lib
from ._lib import check, current_stream, ptr

EPI_STORE_F32, EPI_BIAS_BF16, EPI_BIAS_RELU_BF16, EPI_MASK_BF16 = 0, 1, 2, 3
OPT_SGD, OPT_MOMENTUM, OPT_ADAM = 0, 1, 2
MAX_SEGMENTS = 16

P, I, F = c_void_p, c_int, c_float
_lib.register("dtm_conv1_fwd", [P, I, I, P, P, I, I, P, P, P, P])
_lib.register("dtm_conv1_wgrad", [P, I, I, P, P, I, P, P, P])
_lib.register("dtm_conv2_fwd", [P, P, P, P, I, I, P, P, P, P])
_lib.register("dtm_conv2_dgrad", [P, P, P, P, P, I, I, P, P])
_lib.register("dtm_conv2_pack_weights", [P, P, P, P])
_lib.register("dtm_conv2_wgrad", [P, P, I, P, P])
_lib.register("dtm_conv_wgrad_fused", [P, P, P, I, P, P, I, P, P, P, P])
_lib.register("dtm_unpool2", [P, P, P, I, I, P, P, P])
_lib.register("dtm_colsum_bf16", [P, I, I, P, P])
_lib.register("dtm_set_pdl_mode", [I])
_lib.register("dtm_gemm_mask_unpool", [P, P, P, P, P, P, I, I, P])
_lib.register("dtm_head_forward", [P, I, I, P, P, P, F, F, P, P, P, P, F, P, P])
_lib.register("dtm_head_backward", [P, I, I, P, P, F, P, P, P, P, I, P])
_lib.register("dtm_fused_update", [P, P, P, P, P, P, P, P, F, P])
_lib.register("dtm_fused_update_range", [P, P, P, P, P, P, P, P, F, I, I, I, I, P])
_lib.register("dtm_set_update_l2_window", [P, ctypes.c_size_t, P, P])
_lib.register("dtm_state_init", [P, c_longlong, F, F, P])
_lib.register("dtm_tensor_stats_scratch_bytes", [ctypes.c_size_t], ctypes.c_size_t)
_lib.register("dtm_tensor_stats", [P, ctypes.c_size_t, I, P, P, ctypes.c_size_t, P])
_lib.register("dtm_histogram", [P, ctypes.c_size_t, I, P, I, P, P])


class TrainState(ctypes.Structure):
    """Mirror of ``dtm_train_state`` (include/dtm_b200.h)."""

    _fields_ = [
        ("global_step", ctypes.c_longlong),
        ("beta1_power", c_float), ("beta2_power", c_float),
        ("loss_sum", c_float), ("correct", c_float),
        ("nan_flag", ctypes.c_uint), ("ticket", ctypes.c_uint),
        ("last_xent", c_float), ("last_total", c_float), ("last_acc", c_float), ("last_lr", c_float),
        ("wl_sum", c_float * MAX_SEGMENTS),
        ("last_wl", c_float * MAX_SEGMENTS),
        ("loss_avg", c_float * (MAX_SEGMENTS + 2)),
        ("work", ctypes.c_uint), ("reserved0", ctypes.c_uint),
    ]


class UpdateConfig(ctypes.Structure):
    """Mirror of ``dtm_update_config``."""

    _fields_ = [
        ("optimizer", c_int), ("n_segments", c_int), ("total", c_int), ("reserved", c_int),
        ("decay_steps", ctypes.c_longlong),
        ("base_lr", c_float), ("decay_rate", c_float), ("momentum", c_float), ("beta1", c_float), ("beta2", c_float),
        ("epsilon", c_float), ("ema_decay", c_float),
        ("seg_offset", c_int * (MAX_SEGMENTS + 1)),
        ("seg_wd", c_float * MAX_SEGMENTS),
        ("seg_zero_grad", c_int * MAX_SEGMENTS),
        ("seg_gather_bf16", c_int * MAX_SEGMENTS),
        ("pack_seg", c_int), ("reserved2", c_int), ("pack_fwd", c_void_p), ("pack_dgrad", c_void_p),
    ]


STATE_BYTES = ctypes.sizeof(TrainState)


def _L():
    return _lib.load()


def gemm_bf16(a, a_mn_major, b, b_mn_major, d, M, N, K, *, epilogue=EPI_STORE_F32, bias=None, ref=None, alpha=1.0,
              keep_prob=1.0, seed=0, step_ptr=None, split_k=1, ldd=None):
    check(_L().dtm_gemm_bf16(ptr(a), int(a_mn_major), ptr(b), int(b_mn_major), ptr(d), M, N, K, ldd or N, epilogue,
                             ptr(bias), ptr(ref), float(alpha), float(keep_prob), c_uint64(seed),
                             c_void_p(step_ptr) if isinstance(step_ptr, int) else ptr(step_ptr), split_k,
                             current_stream()), "gemm_bf16")


def gemm_mask_unpool(dy, w_bf16, p2, argmax, dp2, dz2_planes, *, batch, hidden):
    """dP2 GEMM of the first fc layer + ReLU mask + max-pool backward into conv2's plane layout (``dtm_gemm_mask_unpool``)."""
    check(_L().dtm_gemm_mask_unpool(ptr(dy), ptr(w_bf16), ptr(p2), ptr(argmax), ptr(dp2), ptr(dz2_planes), int(batch), int(hidden),
                                    current_stream()), "gemm_mask_unpool")


def conv1_fwd(x, w, bias, p1_planes, argmax, pool_raw=None, *, channels=1, lrn=False):
    batch = x.shape[0]
    check(_L().dtm_conv1_fwd(ptr(x), int(x.dtype == torch.float32), channels, ptr(w), ptr(bias), batch, int(lrn),
                             ptr(p1_planes), ptr(argmax), ptr(pool_raw), current_stream()), "conv1_fwd")


def conv1_wgrad(x, grad_pooled, argmax, dw, db, *, channels=1):
    batch = x.shape[0]
    check(_L().dtm_conv1_wgrad(ptr(x), int(x.dtype == torch.float32), channels, ptr(grad_pooled), ptr(argmax), batch,
                               ptr(dw), ptr(db), current_stream()), "conv1_wgrad")


def conv2_fwd(p1_planes, w2_bf16, bias, p2, argmax, act2=None, *, batch, lrn=False, packed=None):
    check(_L().dtm_conv2_fwd(ptr(p1_planes), ptr(w2_bf16), ptr(packed), ptr(bias), batch, int(lrn), ptr(p2), ptr(argmax),
                             ptr(act2), current_stream()), "conv2_fwd")


def conv2_pack_weights(w2_bf16, packed_fwd, packed_dgrad):
    """Re-pack the bf16 conv2 filter into the two 102400-byte operand images (see include/dtm_b200.h)."""
    check(_L().dtm_conv2_pack_weights(ptr(w2_bf16), ptr(packed_fwd), ptr(packed_dgrad), current_stream()), "conv2_pack_weights")


def conv2_dgrad(dz2_planes, w2_bf16, p1_planes, pool_raw, dpool1, *, batch, lrn=False, packed=None):
    check(_L().dtm_conv2_dgrad(ptr(dz2_planes), ptr(w2_bf16), ptr(packed), ptr(p1_planes), ptr(pool_raw), batch, int(lrn),
                               ptr(dpool1), current_stream()), "conv2_dgrad")


def conv2_wgrad(p1_planes, dz2_planes, dw2, *, batch):
    check(_L().dtm_conv2_wgrad(ptr(p1_planes), ptr(dz2_planes), batch, ptr(dw2), current_stream()), "conv2_wgrad")


def conv_wgrad_fused(p1_planes, dz2_planes, x_u8, grad_pooled, argmax, dw2, dw1, db1, *, batch, channels=1):
    """conv2 and conv1 (uint8 input, 1 or 3 channels) weight gradients in one tcgen05 kernel (``dtm_conv_wgrad_fused``)."""
    if x_u8.dtype != torch.uint8:
        raise TypeError("conv_wgrad_fused: uint8 pixels only (use conv1_wgrad + conv2_wgrad for float input)")
    check(_L().dtm_conv_wgrad_fused(ptr(p1_planes), ptr(dz2_planes), ptr(x_u8), int(channels), ptr(grad_pooled), ptr(argmax), int(batch),
                                    ptr(dw2), ptr(dw1), ptr(db1), current_stream()), "conv_wgrad_fused")


def unpool2(dp2, argmax, act2, dz2_planes, db2, *, batch, lrn=False):
    check(_L().dtm_unpool2(ptr(dp2), ptr(argmax), ptr(act2), batch, int(lrn), ptr(dz2_planes), ptr(db2), current_stream()),
          "unpool2")


def colsum_bf16(x, out):
    M, N = x.shape
    check(_L().dtm_colsum_bf16(ptr(x), M, N, ptr(out), current_stream()), "colsum_bf16")


class pdl_mode:
    """``with pdl_mode(2):`` — programmatic-dependent-launch mode for the launches inside the block (0 off, 1 strict,
    2 early weights; see ``dtm_set_pdl_mode``).  The ``DTM_PDL`` environment variable, when set, wins: it is the A/B
    switch of the tools."""

    def __init__(self, mode: int):
        self.mode = mode
        self.prev = None

    def __enter__(self):
        import os
        if os.environ.get("DTM_PDL") is None:
            self.prev = _L().dtm_set_pdl_mode(int(self.mode))
        return self

    def __exit__(self, *exc):
        if self.prev is not None:
            _L().dtm_set_pdl_mode(int(self.prev))
        return False


def head_forward(h, w, bias, labels, *, grad_scale, loss_scale, logits=None, probs=None, dlogits=None, state=None, dh=None,
                 inv_keep=1.0):
    """fc10 + softmax cross-entropy (+ dlogits, + dH when ``dh`` is given: see ``dtm_head_forward``)."""
    batch, hidden = h.shape
    check(_L().dtm_head_forward(ptr(h), batch, hidden, ptr(w), ptr(bias), ptr(labels), float(grad_scale), float(loss_scale),
                                ptr(logits), ptr(probs), ptr(dlogits), ptr(dh), float(inv_keep), ptr(state), current_stream()),
          "head_forward")


def head_backward(h, w, dlogits, dh, dw, db_out, db_hidden, *, inv_keep=1.0, dh_ready=False):
    batch, hidden = h.shape
    check(_L().dtm_head_backward(ptr(h), batch, hidden, ptr(w), ptr(dlogits), float(inv_keep), ptr(dh), ptr(dw), ptr(db_out),
                                 ptr(db_hidden), int(dh_ready), current_stream()), "head_backward")


def fused_update(cfg: UpdateConfig, state, w, grad, slot0, slot1, ema, w_bf16, *, grad_scale=1.0):
    check(_L().dtm_fused_update(ctypes.byref(cfg), ptr(state), ptr(w), ptr(grad), ptr(slot0), ptr(slot1), ptr(ema), ptr(w_bf16),
                                float(grad_scale), current_stream()), "fused_update")


def fused_update_range(cfg: UpdateConfig, state, w, grad, slot0, slot1, ema, w_bf16, lo, hi, *, grad_scale=1.0, finish=True,
                       max_ctas=0):
    check(_L().dtm_fused_update_range(ctypes.byref(cfg), ptr(state), ptr(w), ptr(grad), ptr(slot0), ptr(slot1), ptr(ema),
                                      ptr(w_bf16), float(grad_scale), int(lo), int(hi), int(finish), int(max_ctas),
                                      current_stream()),
          "fused_update_range")


def set_update_l2_window(blob: torch.Tensor | None) -> dict:
    """Launch the optimizer update kernels of the current device with a persisting-L2 window over ``blob`` (the contiguous
    optimizer state); ``None`` switches it off.  Returns the carve-out and the window limit of the device."""
    carve, win = ctypes.c_size_t(0), ctypes.c_size_t(0)
    nbytes = 0 if blob is None else blob.numel() * blob.element_size()
    check(_L().dtm_set_update_l2_window(ptr(blob), nbytes, ctypes.byref(carve), ctypes.byref(win)), "set_update_l2_window")
    return {"bytes": nbytes, "carveout_bytes": carve.value, "max_window_bytes": win.value}


def state_init(state, global_step=0, beta1=0.9, beta2=0.999):
    check(_L().dtm_state_init(ptr(state), int(global_step), float(beta1), float(beta2), current_stream()), "state_init")


def new_state(device) -> torch.Tensor:
    """Allocate a device-resident ``dtm_train_state`` (as a uint8 tensor) and zero-initialise it."""
    st = torch.zeros(STATE_BYTES, dtype=torch.uint8, device=device)
    return st


def read_state(state: torch.Tensor) -> TrainState:
    """Synchronising device->host read of the training state."""
    raw = bytes(state.cpu().numpy().tobytes())
    return TrainState.from_buffer_copy(raw)


def tensor_stats_scratch_bytes(n: int) -> int:
    return int(_L().dtm_tensor_stats_scratch_bytes(n))


def tensor_stats(x: torch.Tensor, out6: torch.Tensor, scratch: torch.Tensor) -> None:
    """min / max / sum / sumsq / count / L2-norm of a float32 or bfloat16 CUDA tensor into ``out6`` (6 floats)."""
    if x.dtype not in (torch.float32, torch.bfloat16):
        raise TypeError("tensor_stats: float32 or bfloat16 only")
    x = x.contiguous()
    check(_L().dtm_tensor_stats(ptr(x), x.numel(), int(x.dtype == torch.bfloat16), ptr(out6), ptr(scratch),
                                scratch.numel() * scratch.element_size(), current_stream()), "tensor_stats")


def histogram(x: torch.Tensor, limits: torch.Tensor, counts: torch.Tensor) -> None:
    x = x.contiguous()
    check(_L().dtm_histogram(ptr(x), x.numel(), int(x.dtype == torch.bfloat16), ptr(limits), limits.numel(), ptr(counts),
                             current_stream()), "histogram")
