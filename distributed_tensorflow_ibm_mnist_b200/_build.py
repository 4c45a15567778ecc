"""In-tree builder for libdtm_b200.so (the sm_100a kernel library).

The library is plain CUDA C++ with a C ABI (``include/dtm_b200.h``); it is compiled directly with nvcc for
``sm_100a`` — no torch headers, so a full rebuild takes seconds — and loaded with ctypes by
``distributed_tensorflow_ibm_mnist_b200.ops._lib``.  The resulting ``.so`` lives inside the package directory so
that it travels with a snapshot of the repository and is visible to whoever audits which native code was loaded.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent
CSRC = REPO / "csrc"
OUT_DIR = Path(__file__).resolve().parent / "_native"
LIB_PATH = OUT_DIR / "libdtm_b200.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
    "-DDTM_BUILD",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found; libdtm_b200 cannot be built")


def sources() -> list[Path]:
    return sorted(p for p in CSRC.rglob("*") if p.suffix in (".cu", ".cpp"))


def _headers() -> list[Path]:
    return sorted(list(CSRC.rglob("*.cuh")) + list(CSRC.rglob("*.h")) + list((REPO / "include").glob("*.h")))


def _digest(paths: list[Path]) -> str:
    h = hashlib.sha256()
    for p in paths:
        h.update(str(p.relative_to(REPO)).encode())
        h.update(p.read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every source under csrc/ for sm_100a and link libdtm_b200.so. Returns the library path."""
    OUT_DIR.mkdir(parents=True, exist_ok=True)
    obj_dir = OUT_DIR / "obj"
    obj_dir.mkdir(exist_ok=True)
    srcs = sources()
    hdr_digest = _digest(_headers())
    nvcc = _nvcc()

    def compile_one(src: Path) -> tuple[Path, str]:
        obj = obj_dir / (str(src.relative_to(CSRC)).replace("/", "__") + ".o")
        stamp = obj.with_suffix(".stamp")
        want = _digest([src]) + hdr_digest
        log = obj.with_suffix(".log")            # ptxas -v output of this object: ptxas.log stays complete when only
        if not force and obj.exists() and stamp.exists() and stamp.read_text() == want:     # some sources are recompiled
            return obj, log.read_text() if log.exists() else ""
        cmd = [nvcc, *NVCC_FLAGS, "-I", str(REPO / "include"), "-c", str(src), "-o", str(obj)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        stamp.write_text(want)
        log.write_text(r.stderr)
        return obj, r.stderr

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 2)) as ex:
        results = list(ex.map(compile_one, srcs))
    objs = [o for o, _ in results]
    log = "".join(f"== {s.relative_to(REPO)}\n{l}" for s, (_, l) in zip(srcs, results) if l)
    if log:
        (OUT_DIR / "ptxas.log").write_text(log)
        if verbose:
            print(log)
    link_stamp = OUT_DIR / "link.stamp"
    want_link = _digest(srcs) + hdr_digest
    if force or not LIB_PATH.exists() or not link_stamp.exists() or link_stamp.read_text() != want_link:
        cmd = [nvcc, "-shared", "-o", str(LIB_PATH), *map(str, objs), "-cudart", "static", "-lpthread", "-ldl", "-lrt"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
        link_stamp.write_text(want_link)
    return LIB_PATH


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(path)
