#!/usr/bin/env python
"""Hardware-truth probes for tcgen05 descriptors (run on a B200: ``python tools/umma_probe.py --all``).

Each experiment builds a shared-memory byte image + raw UMMA descriptors in numpy, lets
``dtm_umma_probe`` execute them, and compares the TMEM dump with a float32 matmul.  Experiments run in
separate processes so a trapped kernel (watchdog) cannot poison the others.  Results: gpurun_out/probe.json.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))

SW_NONE, SW_128, SW_64, SW_32 = 0, 2, 4, 6


def bf16_round(x: np.ndarray) -> np.ndarray:
    import torch

    return torch.from_numpy(x.astype(np.float32)).to(torch.bfloat16).to(torch.float32).numpy()


def bf16_bits(x: np.ndarray) -> np.ndarray:
    import torch

    return torch.from_numpy(x.astype(np.float32)).to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)


def smem_desc(addr: int, lbo: int, sbo: int, swizzle: int) -> int:
    assert addr % 16 == 0 and lbo % 16 == 0 and sbo % 16 == 0
    d = (addr >> 4) & 0x3FFF
    d |= ((lbo >> 4) & 0x3FFF) << 16
    d |= ((sbo >> 4) & 0x3FFF) << 32
    d |= 1 << 46
    d |= (swizzle & 7) << 61
    return d


def idesc(M: int, N: int, a_mn: bool, b_mn: bool) -> int:
    return (1 << 4) | (1 << 7) | (1 << 10) | (int(a_mn) << 15) | (int(b_mn) << 16) | ((N >> 3) << 17) | ((M >> 4) << 24)


def run_probe(image: np.ndarray, ops: list[tuple[int, int, int, int, int]], ncols: int) -> np.ndarray:
    import torch
    from distributed_tensorflow_ibm_mnist_b200.ops import _lib

    lib = _lib.load()
    img = torch.from_numpy(image.view(np.uint8).copy()).cuda()
    rec = np.zeros(len(ops), dtype=[("a", "<u8"), ("b", "<u8"), ("i", "<u4"), ("col", "<u4"), ("acc", "<u4"), ("pad", "<u4")])
    for k, (da, db, ide, col, acc) in enumerate(ops):
        rec[k] = (da, db, ide, col, acc, 0)
    ops_t = torch.from_numpy(rec.view(np.uint8).copy()).cuda()
    dump = torch.full((128, ncols), float("nan"), device="cuda", dtype=torch.float32)
    _lib.check(lib.dtm_umma_probe(_lib.ptr(img), img.numel(), _lib.ptr(ops_t), len(ops), _lib.ptr(dump), ncols,
                                  _lib.current_stream()), "umma_probe")
    torch.cuda.synchronize()
    return dump.cpu().numpy()


def put16(image: np.ndarray, byte_off: int, vals8: np.ndarray) -> None:
    """Store 8 bf16 values (one 16-byte chunk) at byte offset."""
    assert byte_off % 16 == 0
    image[byte_off // 2: byte_off // 2 + 8] = vals8


def report(name: str, got: np.ndarray, want: np.ndarray, extra: dict | None = None) -> dict:
    err = float(np.nanmax(np.abs(got - want))) if np.isfinite(got).all() else float("inf")
    scale = float(np.abs(want).max()) + 1e-9
    res = {"exp": name, "max_abs_err": err, "ref_scale": scale, "pass": bool(err <= 2e-2 * scale)}
    if extra:
        res.update(extra)
    return res


# ------------------------------------------------------------------------------------------------
def exp_kmajor_noswz(swap: bool = False) -> dict:
    """K-major, no swizzle: layout [k8][row][8]; LBO = K-direction stride, SBO = 8-row-group stride."""
    rng = np.random.default_rng(1)
    M, N, K = 128, 64, 16
    A = bf16_round(rng.standard_normal((M, K)))
    B = bf16_round(rng.standard_normal((N, K)))
    image = np.zeros(32 * 1024 // 2, dtype=np.uint16)
    a_off, b_off = 0, 8192
    Ab, Bb = bf16_bits(A), bf16_bits(B)
    for k8 in range(2):
        for m in range(M):
            put16(image, a_off + k8 * (M * 16) + m * 16, Ab[m, k8 * 8:(k8 + 1) * 8])
        for n in range(N):
            put16(image, b_off + k8 * (N * 16) + n * 16, Bb[n, k8 * 8:(k8 + 1) * 8])
    a_lbo, a_sbo, b_lbo, b_sbo = M * 16, 128, N * 16, 128
    if swap:
        a_lbo, a_sbo, b_lbo, b_sbo = a_sbo, a_lbo, b_sbo, b_lbo
    ops = [(smem_desc(a_off, a_lbo, a_sbo, SW_NONE), smem_desc(b_off, b_lbo, b_sbo, SW_NONE), idesc(M, N, False, False), 0, 0)]
    got = run_probe(image, ops, 64)
    return report("kmajor_noswz" + ("_swapped" if swap else ""), got[:, :N], A @ B.T)


def exp_conv_shift() -> dict:
    """The zero-copy conv tap: A rows are pixels of a plane layout, tap = re-based start address."""
    rng = np.random.default_rng(2)
    NPX, C, N = 512, 32, 64
    P = bf16_round(rng.standard_normal((NPX, C)))       # [pixel][channel]
    W = bf16_round(rng.standard_normal((N, C)))          # [n][k=channel]
    plane_pitch = NPX * 16
    image = np.zeros(96 * 1024 // 2, dtype=np.uint16)
    Pb, Wb = bf16_bits(P), bf16_bits(W)
    for c8 in range(C // 8):
        for px in range(NPX):
            put16(image, c8 * plane_pitch + px * 16, Pb[px, c8 * 8:(c8 + 1) * 8])
    b_off = (C // 8) * plane_pitch
    for k8 in range(C // 8):
        for n in range(N):
            put16(image, b_off + k8 * (N * 16) + n * 16, Wb[n, k8 * 8:(k8 + 1) * 8])
    shift = 37  # pixels (arbitrary, not a multiple of 8)
    ops = []
    for j in range(C // 16):
        da = smem_desc(shift * 16 + j * 2 * plane_pitch, plane_pitch, 128, SW_NONE)
        db = smem_desc(b_off + j * 2 * (N * 16), N * 16, 128, SW_NONE)
        ops.append((da, db, idesc(128, N, False, False), 0, 1 if j else 0))
    got = run_probe(image, ops, 64)
    return report("conv_shift", got[:, :N], P[shift:shift + 128] @ W.T)


def exp_mnmajor_noswz(M: int) -> dict:
    """MN-major, no swizzle, from the plane layout: M = channels (planes), K = pixels."""
    rng = np.random.default_rng(3)
    NPX, N, K = 64, 32, 16
    CA = M
    PA = bf16_round(rng.standard_normal((NPX, CA)))      # [pixel][channel]  -> A^T
    PB = bf16_round(rng.standard_normal((NPX, N)))
    pitch = NPX * 16
    image = np.zeros(64 * 1024 // 2, dtype=np.uint16)
    for c8 in range(CA // 8):
        for px in range(NPX):
            put16(image, c8 * pitch + px * 16, bf16_bits(PA[px, c8 * 8:(c8 + 1) * 8]))
    b_off = (CA // 8) * pitch
    for c8 in range(N // 8):
        for px in range(NPX):
            put16(image, b_off + c8 * pitch + px * 16, bf16_bits(PB[px, c8 * 8:(c8 + 1) * 8]))
    shift = 5
    # MN-major no swizzle: SBO = stride between 8-element MN groups (planes), LBO = stride between 8-k groups
    da = smem_desc(0, 128, pitch, SW_NONE)
    db = smem_desc(b_off + shift * 16, 128, pitch, SW_NONE)
    ops = [(da, db, idesc(M, N, True, True), 0, 0)]
    got = run_probe(image, ops, 32)
    want = PA[:K].T @ PB[shift:shift + K]                # [M][N]
    res = {}
    if M == 128:
        return report("mnmajor_noswz_m128", got[:, :N], want)
    # M = 64: discover which TMEM lanes hold which rows
    lane_of_row = []
    for r in range(64):
        d = np.abs(got[:, :N] - want[r][None, :]).max(axis=1)
        lane_of_row.append(int(np.argmin(d)) if np.isfinite(d).any() else -1)
    ok = all(np.abs(got[l, :N] - want[r]).max() < 2e-2 * (np.abs(want).max() + 1e-9) for r, l in enumerate(lane_of_row))
    return {"exp": "mnmajor_noswz_m64", "pass": bool(ok), "lane_of_row": lane_of_row}


def swz128(row: int, chunk: int) -> int:
    return chunk ^ (row & 7)


def exp_kmajor_sw128() -> dict:
    rng = np.random.default_rng(4)
    M, N, K = 128, 64, 64
    A = bf16_round(rng.standard_normal((M, K)))
    B = bf16_round(rng.standard_normal((N, K)))
    image = np.zeros(32 * 1024 // 2, dtype=np.uint16)
    a_off, b_off = 0, 16384
    Ab, Bb = bf16_bits(A), bf16_bits(B)
    for m in range(M):
        for c in range(8):
            put16(image, a_off + m * 128 + swz128(m, c) * 16, Ab[m, c * 8:(c + 1) * 8])
    for n in range(N):
        for c in range(8):
            put16(image, b_off + n * 128 + swz128(n, c) * 16, Bb[n, c * 8:(c + 1) * 8])
    ops = []
    for j in range(4):
        ops.append((smem_desc(a_off + j * 32, 0, 1024, SW_128), smem_desc(b_off + j * 32, 0, 1024, SW_128),
                    idesc(M, N, False, False), 0, 1 if j else 0))
    got = run_probe(image, ops, 64)
    return report("kmajor_sw128", got[:, :N], A @ B.T)


def exp_mnmajor_sw128() -> dict:
    rng = np.random.default_rng(5)
    M, N, K = 128, 64, 64
    At = bf16_round(rng.standard_normal((K, M)))   # [k][m]
    Bt = bf16_round(rng.standard_normal((K, N)))   # [k][n]
    image = np.zeros(32 * 1024 // 2, dtype=np.uint16)
    a_off, b_off = 0, 16384
    Ab, Bb = bf16_bits(At), bf16_bits(Bt)
    for atom in range(2):
        for k in range(K):
            for c in range(8):
                put16(image, a_off + atom * (K * 128) + k * 128 + swz128(k, c) * 16,
                      Ab[k, atom * 64 + c * 8: atom * 64 + (c + 1) * 8])
    for k in range(K):
        for c in range(8):
            put16(image, b_off + k * 128 + swz128(k, c) * 16, Bb[k, c * 8:(c + 1) * 8])
    ops = []
    for j in range(4):
        ops.append((smem_desc(a_off + j * 2048, K * 128, 1024, SW_128), smem_desc(b_off + j * 2048, K * 128, 1024, SW_128),
                    idesc(M, N, True, True), 0, 1 if j else 0))
    got = run_probe(image, ops, 64)
    return report("mnmajor_sw128", got[:, :N], At.T @ Bt)


def exp_sw64_shift_stack() -> dict:
    """MN-major SW64 with overlapping atoms: M = 4 pixel shifts x 32 channels from a dense [px][32ch] layout."""
    rng = np.random.default_rng(6)
    NPX, C, N, K = 128, 32, 64, 16
    P = bf16_round(rng.standard_normal((NPX, C)))
    Q = bf16_round(rng.standard_normal((NPX, N)))
    image = np.zeros(64 * 1024 // 2, dtype=np.uint16)
    Pb = bf16_bits(P)
    for p in range(NPX):
        for c in range(4):
            put16(image, p * 64 + ((c ^ ((p >> 1) & 3)) * 16), Pb[p, c * 8:(c + 1) * 8])
    b_off = 16384
    pitch = NPX * 16
    for c8 in range(N // 8):
        for px in range(NPX):
            put16(image, b_off + c8 * pitch + px * 16, bf16_bits(Q[px, c8 * 8:(c8 + 1) * 8]))
    base_px = 8
    da = smem_desc(base_px * 64, 64, 512, SW_64)       # LBO: next MN atom = next pixel; SBO: 8 k-rows
    db = smem_desc(b_off + base_px * 16, 128, pitch, SW_NONE)
    ops = [(da, db, idesc(128, N, True, True), 0, 0)]
    got = run_probe(image, ops, 64)
    want = np.zeros((128, N), dtype=np.float32)
    for j in range(4):
        want[j * 32:(j + 1) * 32] = P[base_px + j: base_px + j + K].T @ Q[base_px: base_px + K]
    return report("sw64_shift_stack", got[:, :N], want)


def exp_gemm(a_mn: int, b_mn: int, M: int, N: int, K: int, epi: int = 0, split: int = 1) -> dict:
    import torch
    from distributed_tensorflow_ibm_mnist_b200.ops import _lib

    lib = _lib.load()
    g = torch.Generator(device="cpu").manual_seed(7)
    A = torch.randn(M, K, generator=g).to(torch.bfloat16)
    B = torch.randn(K, N, generator=g).to(torch.bfloat16)
    want = A.float() @ B.float()
    a_dev = (A.t().contiguous() if a_mn else A.contiguous()).cuda()
    b_dev = (B.contiguous() if b_mn else B.t().contiguous()).cuda()
    if epi == 0:
        d = torch.zeros(M, N, device="cuda", dtype=torch.float32)
    else:
        d = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16)
    bias = torch.randn(N, generator=g).cuda() if epi in (1, 2) else None
    _lib.check(lib.dtm_gemm_bf16(_lib.ptr(a_dev), a_mn, _lib.ptr(b_dev), b_mn, _lib.ptr(d), M, N, K, N, epi,
                                 _lib.ptr(bias), None, 1.0, 1.0, 0, None, split, _lib.current_stream()), "gemm")
    torch.cuda.synchronize()
    if epi == 1:
        want = want + bias.cpu()
    if epi == 2:
        want = torch.relu(want + bias.cpu())
    got = d.float().cpu()
    err = (got - want).abs().max().item()
    scale = want.abs().max().item()
    tol = 2e-2 if epi else 1e-3
    return {"exp": f"gemm_a{a_mn}b{b_mn}_{M}x{N}x{K}_epi{epi}_s{split}", "max_abs_err": err, "ref_scale": scale,
            "pass": bool(err <= tol * scale)}


EXPERIMENTS = {
    "kmajor_noswz": lambda: exp_kmajor_noswz(False),
    "kmajor_noswz_swapped": lambda: exp_kmajor_noswz(True),
    "conv_shift": exp_conv_shift,
    "mnmajor_noswz_m128": lambda: exp_mnmajor_noswz(128),
    "mnmajor_noswz_m64": lambda: exp_mnmajor_noswz(64),
    "kmajor_sw128": exp_kmajor_sw128,
    "mnmajor_sw128": exp_mnmajor_sw128,
    "sw64_shift_stack": exp_sw64_shift_stack,
    "gemm_kk": lambda: exp_gemm(0, 0, 512, 3136, 1024),
    "gemm_kmn": lambda: exp_gemm(0, 1, 512, 1024, 3136, epi=2),
    "gemm_mnmn": lambda: exp_gemm(1, 1, 3136, 1024, 512),
    "gemm_mnk": lambda: exp_gemm(1, 0, 256, 192, 1024),
    "gemm_small": lambda: exp_gemm(0, 1, 128, 64, 64),
    "gemm_splitk": lambda: exp_gemm(0, 1, 512, 1024, 3136, epi=0, split=4),
    "gemm_ragged": lambda: exp_gemm(0, 1, 200, 192, 1000, epi=1),
}


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--exp")
    ap.add_argument("--all", action="store_true")
    ap.add_argument("--out", default=str(REPO / "gpurun_out" / "probe.json"))
    args = ap.parse_args()
    if args.exp:
        res = EXPERIMENTS[args.exp]()
        print("PROBE_RESULT " + json.dumps(res))
        return
    results = []
    for name in EXPERIMENTS:
        try:
            r = subprocess.run([sys.executable, __file__, "--exp", name], capture_output=True, text=True, timeout=180)
            line = [l for l in r.stdout.splitlines() if l.startswith("PROBE_RESULT ")]
            if line:
                results.append(json.loads(line[-1][len("PROBE_RESULT "):]))
            else:
                results.append({"exp": name, "pass": False, "rc": r.returncode, "stdout": r.stdout[-600:], "stderr": r.stderr[-1200:]})
        except subprocess.TimeoutExpired:
            results.append({"exp": name, "pass": False, "error": "timeout"})
        print(json.dumps(results[-1]), flush=True)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    Path(args.out).write_text(json.dumps(results, indent=1))


if __name__ == "__main__":
    main()
