"""NumPy mirror of csrc/common/philox.cuh: the dropout mask as a pure function of (seed, step, element index).

The GPU kernels never store a dropout mask — forward and backward regenerate it from the counter — so the oracle
needs the identical generator to reproduce a training step bit-for-bit on the host.
"""
from __future__ import annotations

import numpy as np

_M0, _M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
_W0, _W1 = np.uint32(0x9E3779B9), np.uint32(0xBB67AE85)
_LO = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised Philox4x32-10. Counters are uint32 arrays, keys uint32 scalars. Returns four uint32 arrays."""
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint32) for c in (c0, c1, c2, c3))
    k0, k1 = np.uint32(k0), np.uint32(k1)
    with np.errstate(over="ignore"):
        for _ in range(10):
            p0 = _M0 * c0.astype(np.uint64)
            p1 = _M1 * c2.astype(np.uint64)
            n0 = (p1 >> np.uint64(32)).astype(np.uint32) ^ c1 ^ k0
            n1 = (p1 & _LO).astype(np.uint32)
            n2 = (p0 >> np.uint64(32)).astype(np.uint32) ^ c3 ^ k1
            n3 = (p0 & _LO).astype(np.uint32)
            c0, c1, c2, c3 = n0, n1, n2, n3
            k0 = np.uint32((int(k0) + int(_W0)) & 0xFFFFFFFF)
            k1 = np.uint32((int(k1) + int(_W1)) & 0xFFFFFFFF)
    return c0, c1, c2, c3


def dropout_keep_mask(rows: int, cols: int, *, seed: int, step: int, keep_prob: float) -> np.ndarray:
    """Boolean [rows, cols] keep mask identical to the GEMM epilogue's (cols must be a multiple of 4)."""
    assert cols % 4 == 0
    n4 = rows * cols // 4
    e4 = np.arange(n4, dtype=np.uint64)
    c0 = (e4 & _LO).astype(np.uint32)
    c1 = (e4 >> np.uint64(32)).astype(np.uint32)
    c2 = np.full(n4, step & 0xFFFFFFFF, dtype=np.uint32)
    c3 = np.full(n4, (step >> 32) & 0xFFFFFFFF, dtype=np.uint32)
    r = philox4x32_10(c0, c1, c2, c3, seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    thresh = np.uint32(int(np.float32(keep_prob) * np.float32(16777216.0)))
    bits = np.stack(r, axis=1).reshape(-1)
    return ((bits >> np.uint32(8)) < thresh).reshape(rows, cols)
