"""Host/torch helpers for the *plane layout* (csrc/conv/plane_layout.cuh) used by the tensor-core conv kernels.

Per image: C/8 planes; a plane is a 16x16 pixel grid with 8 bf16 channels per pixel.  Plane row r holds image row
r-2 (rows 0-1 zero), plane column c holds image column c (columns 14-15 zero).  These converters are used by the
tests (to build inputs / decode outputs) and by tools; the training path itself never converts — conv1 writes the
layout directly and conv2 consumes it with re-based UMMA descriptors.
"""
from __future__ import annotations

import torch

GRID = 16
ROW_OFF = 2
IMG = 14


def plane_bytes(channels: int) -> int:
    return (channels // 8) * GRID * GRID * 16


def alloc_planes(batch: int, channels: int, device) -> torch.Tensor:
    """Zero-initialised plane-layout buffer (zero borders are part of the format and are never rewritten)."""
    return torch.zeros(batch, plane_bytes(channels), dtype=torch.uint8, device=device)


def nhwc_to_planes(x: torch.Tensor) -> torch.Tensor:
    """[B,14,14,C] (any float dtype) -> uint8 [B, C/8*4096] plane layout in bf16."""
    B, H, W, C = x.shape
    assert H == IMG and W == IMG and C % 8 == 0
    grid = torch.zeros(B, C // 8, GRID, GRID, 8, dtype=torch.bfloat16, device=x.device)
    xb = x.to(torch.bfloat16).view(B, H, W, C // 8, 8).permute(0, 3, 1, 2, 4)
    grid[:, :, ROW_OFF:ROW_OFF + H, :W, :] = xb
    return grid.contiguous().view(torch.uint8).view(B, -1)


def planes_to_nhwc(p: torch.Tensor, channels: int) -> torch.Tensor:
    """uint8 plane layout -> bf16 [B,14,14,C]."""
    B = p.shape[0]
    grid = p.contiguous().view(torch.bfloat16).view(B, channels // 8, GRID, GRID, 8)
    x = grid[:, :, ROW_OFF:ROW_OFF + IMG, :IMG, :].permute(0, 2, 3, 1, 4).reshape(B, IMG, IMG, channels)
    return x.contiguous()


def planes_border_is_zero(p: torch.Tensor, channels: int) -> bool:
    B = p.shape[0]
    grid = p.contiguous().view(torch.bfloat16).view(B, channels // 8, GRID, GRID, 8).float()
    return bool((grid[:, :, :ROW_OFF].abs().sum() == 0) and (grid[:, :, :, IMG:].abs().sum() == 0))
