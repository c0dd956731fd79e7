"""Host-side view of the padded planar-chunk activation layout (csrc/layout.cuh).

Used by tests and debugging tools to move between the reference's NCHW fp32 tensors
(reference model.py:18-20) and the bf16 planes the sm_100a kernels read and write.
Pure index arithmetic, works on CPU and CUDA tensors.
"""
from __future__ import annotations

import torch

TILE_M = 128
HALO = 32
GUARD = 64


def tiles_for(batch: int, H: int, W: int) -> int:
    return (batch * (H + 2) * (W + 2) + TILE_M - 1) // TILE_M


def plane_rows(batch_cap: int, H: int, W: int) -> int:
    return GUARD + tiles_for(batch_cap, H, W) * TILE_M + GUARD


def row_index(batch: int, H: int, W: int, device=None) -> torch.Tensor:
    """int64 [batch, H, W]: plane row of every real pixel."""
    PW, PHW = W + 2, (H + 2) * (W + 2)
    b = torch.arange(batch, device=device).view(-1, 1, 1)
    h = torch.arange(H, device=device).view(1, -1, 1)
    w = torch.arange(W, device=device).view(1, 1, -1)
    return GUARD + b * PHW + (h + 1) * PW + (w + 1)


def nchw_to_planes(x: torch.Tensor, batch_cap: int | None = None) -> torch.Tensor:
    """[B,C,H,W] float -> zero-padded bf16 planes [C/8, rows, 8]."""
    B, C, H, W = x.shape
    assert C % 8 == 0
    rows = plane_rows(batch_cap or B, H, W)
    planes = torch.zeros(C // 8, rows, 8, dtype=torch.bfloat16, device=x.device)
    idx = row_index(B, H, W, x.device).reshape(-1)
    # [B,C,H,W] -> [C/8, B*H*W, 8]
    v = x.reshape(B, C // 8, 8, H, W).permute(1, 0, 3, 4, 2).reshape(C // 8, B * H * W, 8)
    planes[:, idx, :] = v.to(torch.bfloat16)
    return planes


def planes_to_nchw(planes: torch.Tensor, B: int, C: int, H: int, W: int) -> torch.Tensor:
    """bf16 planes -> [B,C,H,W] float32 (interior pixels only)."""
    idx = row_index(B, H, W, planes.device).reshape(-1)
    v = planes[:, idx, :].float().reshape(C // 8, B, H, W, 8)
    return v.permute(1, 0, 4, 2, 3).reshape(B, C, H, W).contiguous()


def border_rows_are_zero(planes: torch.Tensor, B: int, H: int, W: int) -> bool:
    """True if every row of the tile range that is not a real pixel holds zeros."""
    ntile_rows = tiles_for(B, H, W) * TILE_M
    keep = torch.ones(planes.shape[1], dtype=torch.bool, device=planes.device)
    keep[row_index(B, H, W, planes.device).reshape(-1)] = False
    keep[:GUARD] = False
    keep[GUARD + ntile_rows:] = False
    return bool((planes[:, keep, :].float() == 0).all())
