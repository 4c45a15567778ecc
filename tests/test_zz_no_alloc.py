"""The allocation-free statistics entry point, exercised through the C ABI.

``dtm_tensor_stats_static`` exists for callers that cannot give the library any device scratch: a diagnostics hook that
runs when the card's memory is completely committed (or while an out-of-memory / NaN post-mortem is written), a
watcher thread that must stay away from cudaMalloc because that call may synchronise the device and stall the training
streams it is watching, or a plain-C program that uses this ABI without PyTorch and has no device allocator.  It keeps
its per-CTA partials in a 16 KiB array inside the library, so it accepts at most 1 044 480 elements.

The test calls it twice in a row — the second call sees whatever the first left in the shared array — and each
result must equal what the regular entry point (caller-provided scratch) returns for the same data, within the
tolerances the other statistics tests use.  It also checks that an over-size request is refused with a message that
names the limit.  This file sorts last on purpose: a run that stops at the first failure has then already executed
every other test, and the test must hold both in a fresh process and after the rest of the suite has used the GPU.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_ELEMS = 1044480


def _lib():
    from distributed_tensorflow_ibm_mnist_b200.ops import _lib, native  # noqa: F401  (native registers the base signatures)

    lib = _lib.load()
    lib.dtm_tensor_stats_static.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    lib.dtm_tensor_stats_static.restype = ctypes.c_int
    return lib


def _both(lib, x):
    """(static-scratch result, caller-scratch result) for one tensor, each through the raw C entry points."""
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    n = x.numel()
    is_bf16 = int(x.dtype == torch.bfloat16)
    out_static = torch.full((6,), float("nan"), device="cuda")
    rc = lib.dtm_tensor_stats_static(ctypes.c_void_p(x.data_ptr()), n, is_bf16, ctypes.c_void_p(out_static.data_ptr()), stream)
    assert rc == 0, lib.dtm_last_error().decode()
    scratch = torch.zeros(int(lib.dtm_tensor_stats_scratch_bytes(n)), dtype=torch.uint8, device="cuda")
    out_ref = torch.full((6,), float("nan"), device="cuda")
    rc = lib.dtm_tensor_stats(ctypes.c_void_p(x.data_ptr()), n, is_bf16, ctypes.c_void_p(out_ref.data_ptr()),
                              ctypes.c_void_p(scratch.data_ptr()), scratch.numel(), stream)
    assert rc == 0, lib.dtm_last_error().decode()
    torch.cuda.synchronize()
    return out_static.cpu().double(), out_ref.cpu().double()


def test_static_scratch_matches_regular_entry_point_twice_in_a_row():
    lib = _lib()
    g = torch.Generator().manual_seed(2024)
    first = (torch.randn(MAX_ELEMS, generator=g) * 0.05).cuda()                      # exactly at the limit: 510 CTAs
    second = (torch.randn(196 * 1024 + 3, generator=g) * 3.0 + 1.0).to(torch.bfloat16).cuda()   # other size, other dtype
    for x in (first, second):
        got, want = _both(lib, x)
        assert torch.equal(got[:2], want[:2])                       # min / max are exact
        assert torch.equal(got[4:5], want[4:5])                     # element count
        assert torch.allclose(got[2:4], want[2:4], rtol=1e-5, atol=1e-6)
        assert torch.allclose(got[5:], want[5:], rtol=1e-5)
        xf = x.double().cpu()
        assert abs(got[3] - (xf * xf).sum()) <= 1e-5 * (xf * xf).sum()   # and both agree with the host value


def test_static_scratch_rejects_oversize_input_with_a_clear_message():
    lib = _lib()
    x = torch.zeros(MAX_ELEMS + 1, device="cuda")
    out = torch.zeros(6, device="cuda")
    rc = lib.dtm_tensor_stats_static(ctypes.c_void_p(x.data_ptr()), x.numel(), 0, ctypes.c_void_p(out.data_ptr()),
                                     ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0
    msg = lib.dtm_last_error().decode()
    assert "1044480" in msg and "too large" in msg
    assert float(out.abs().sum()) == 0.0            # nothing was launched
