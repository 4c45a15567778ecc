"""CRC-32C (Castagnoli) and TensorFlow's masked variant.

Used by every on-disk format the reference relies on through TensorFlow: TFRecord framing
(reference mnist_input.py:27, inference.py:50), tensor-bundle checkpoints (main.py:143, inference.py:86-91)
and event files (main.py:78).  The byte loop runs in libdtm_b200 (SSE4.2 ``crc32`` instruction) when the
native library is present; the pure-Python table version is the always-available fallback and the test oracle.
"""
from __future__ import annotations

import ctypes

_POLY = 0x82F63B78
_TABLE: list[int] = []
for _i in range(256):
    _c = _i
    for _ in range(8):
        _c = (_c >> 1) ^ _POLY if _c & 1 else _c >> 1
    _TABLE.append(_c)

_MASK_DELTA = 0xA282EAD8
_native = None


def _crc32c_py(data: bytes, crc: int = 0) -> int:
    c = crc ^ 0xFFFFFFFF
    tab = _TABLE
    for b in data:
        c = tab[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def _load_native():
    global _native
    if _native is None:
        try:
            from ..ops import _lib

            lib = _lib.load()
            fn = lib.dtm_crc32c
            fn.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_uint32]
            fn.restype = ctypes.c_uint32
            _native = fn
        except Exception:  # library not built: stay on the Python path
            _native = False
    return _native


def crc32c(data, crc: int = 0) -> int:
    """CRC-32C of ``data`` (bytes-like), optionally continuing from a previous value."""
    if not isinstance(data, (bytes, bytearray)):
        data = bytes(data)
    if len(data) >= 256:
        fn = _load_native()
        if fn:
            return int(fn(bytes(data), len(data), crc))
    return _crc32c_py(data, crc)


def mask(crc: int) -> int:
    """TensorFlow/LevelDB masking: rotate right by 15 and add a constant (so CRCs of CRCs stay well mixed)."""
    return (((crc >> 15) | (crc << 17)) + _MASK_DELTA) & 0xFFFFFFFF


def unmask(masked: int) -> int:
    rot = (masked - _MASK_DELTA) & 0xFFFFFFFF
    return ((rot >> 17) | (rot << 15)) & 0xFFFFFFFF


def masked_crc32c(data) -> int:
    return mask(crc32c(data))
