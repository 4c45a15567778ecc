"""ctypes binding to libdtm_b200.so (C ABI in include/dtm_b200.h).

The kernels are the *only* GPU compute path of this framework: if the shared object is missing, or reports a
different ABI than this binding was written for, loading fails loudly instead of falling back to anything else.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_uint32, c_uint64, c_void_p
from pathlib import Path

ABI_VERSION = 11
_LIB = None
_LIB_PATH = Path(__file__).resolve().parent.parent / "_native" / "libdtm_b200.so"


class NativeError(RuntimeError):
    pass


def lib_path() -> Path:
    return _LIB_PATH


def available() -> bool:
    return _LIB_PATH.exists()


def _declare(lib: ctypes.CDLL) -> None:
    P, I, F, U32, U64 = c_void_p, c_int, c_float, c_uint32, c_uint64
    sigs = {
        "dtm_abi_version": ([], I),
        "dtm_last_error": ([], c_char_p),
        "dtm_device_info": ([P, P, P, P], I),
        "dtm_umma_probe": ([P, U32, P, U32, P, U32, P], I),
        "dtm_gemm_bf16": ([P, I, P, I, P, I, I, I, I, I, P, P, F, F, U64, P, I, P], I),
    }
    for name, (args, res) in sigs.items():
        fn = getattr(lib, name, None)
        if fn is None:
            continue
        fn.argtypes = args
        fn.restype = res
    for name, (args, res) in EXTRA_SIGNATURES.items():
        fn = getattr(lib, name, None)
        if fn is None:
            raise NativeError(f"libdtm_b200.so lacks symbol {name}; rebuild with python -m {__package__.split('.')[0]}._build")
        fn.argtypes = args
        fn.restype = res


# Filled by the op modules at import time (name -> (argtypes, restype)); keeps each signature next to its wrapper.
EXTRA_SIGNATURES: dict[str, tuple[list, object]] = {}


def register(name: str, argtypes: list, restype=c_int) -> None:
    EXTRA_SIGNATURES[name] = (argtypes, restype)
    if _LIB is not None:
        fn = getattr(_LIB, name)
        fn.argtypes = argtypes
        fn.restype = restype


def load() -> ctypes.CDLL:
    """Load (once) and return the native library. Builds it first when DTM_AUTOBUILD=1 and it is missing."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not _LIB_PATH.exists():
        if os.environ.get("DTM_AUTOBUILD", "0") == "1":
            from .. import _build

            _build.build()
        else:
            raise NativeError(
                f"{_LIB_PATH} not found. Build it with `python -m distributed_tensorflow_ibm_mnist_b200._build` "
                "(or call __graft_entry__.build())."
            )
    lib = ctypes.CDLL(str(_LIB_PATH), mode=ctypes.RTLD_GLOBAL)
    got = lib.dtm_abi_version()
    if got != ABI_VERSION:
        raise NativeError(f"libdtm_b200.so ABI {got} != binding ABI {ABI_VERSION}; rebuild the library")
    _declare(lib)
    _LIB = lib
    return lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().dtm_last_error()
        raise NativeError(f"{what or 'native call'} failed (rc={rc}): {msg.decode() if msg else '?'}")


def ptr(t) -> c_void_p:
    """Device (or host) pointer of a torch tensor, or NULL for None."""
    if t is None:
        return c_void_p(0)
    return c_void_p(t.data_ptr())


def current_stream() -> c_void_p:
    import torch

    return c_void_p(torch.cuda.current_stream().cuda_stream)
