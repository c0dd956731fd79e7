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


def nchw_to_planes(x: torch.Tensor, batch_cap: int | None = None,
                   dtype: torch.dtype = torch.bfloat16) -> torch.Tensor:
    """[B,C,H,W] float -> zero-padded planes: [C/8, rows, 8] bf16 or, for ``dtype=torch.float32``,
    [C/4, rows, 4] fp32 (16-byte rows either way)."""
    B, C, H, W = x.shape
    ch = 4 if dtype == torch.float32 else 8
    assert C % ch == 0
    rows = plane_rows(batch_cap or B, H, W)
    planes = torch.zeros(C // ch, rows, ch, dtype=dtype, device=x.device)
    idx = row_index(B, H, W, x.device).reshape(-1)
    # [B,C,H,W] -> [C/ch, B*H*W, ch]
    v = x.reshape(B, C // ch, ch, H, W).permute(1, 0, 3, 4, 2).reshape(C // ch, B * H * W, ch)
    planes[:, idx, :] = v.to(dtype)
    return planes


def nchw_to_split_planes(x: torch.Tensor, batch_cap: int | None = None) -> torch.Tensor:
    """[B,C,H,W] fp32 -> split planes [2*C/8, rows, 8] bf16: hi = nearest bf16 of x (planes 0..C/8-1),
    lo = nearest bf16 of (x - hi) (planes C/8..): what the producing kernels store (csrc/layout.cuh)."""
    hi = x.to(torch.bfloat16)
    lo = (x - hi.float()).to(torch.bfloat16)
    return torch.cat([nchw_to_planes(hi.float(), batch_cap), nchw_to_planes(lo.float(), batch_cap)], 0)


def split_value(x: torch.Tensor) -> torch.Tensor:
    """The fp32 value hi + lo that split planes hold for x (a 16-bit significand)."""
    hi = x.to(torch.bfloat16).float()
    return hi + (x - hi).to(torch.bfloat16).float()


def planes_to_nchw(planes: torch.Tensor, B: int, C: int, H: int, W: int) -> torch.Tensor:
    """bf16, split or fp32 planes -> [B,C,H,W] float32 (interior pixels only); split planes are
    recognised by their plane count (2*C/8) and returned as hi + lo."""
    ch = planes.shape[-1]
    if ch == 8 and planes.shape[0] == 2 * (C // 8):
        n = C // 8
        return planes_to_nchw(planes[:n], B, C, H, W) + planes_to_nchw(planes[n:], B, C, H, W)
    idx = row_index(B, H, W, planes.device).reshape(-1)
    v = planes[:, idx, :].float().reshape(C // ch, B, H, W, ch)
    return v.permute(1, 0, 4, 2, 3).reshape(B, C, H, W).contiguous()


def round_tf32(x: torch.Tensor) -> torch.Tensor:
    """Nearest tf32 (10 mantissa bits, ties away from zero) of an fp32 tensor, as fp32 — the
    host-side restatement of ``cvt.rna.tf32.f32`` used by the kernels (csrc/ptx.cuh)."""
    bits = x.contiguous().view(torch.int32)
    out = ((bits + 0x1000) & ~0x1FFF).view(torch.float32)
    return torch.where(torch.isfinite(x), out, x)


def border_rows_are_zero(planes: torch.Tensor, B: int, H: int, W: int) -> bool:
    """True if every row of the tile range that is not a real pixel holds zeros."""
    ntile_rows = tiles_for(B, H, W) * TILE_M
    keep = torch.ones(planes.shape[1], dtype=torch.bool, device=planes.device)
    keep[row_index(B, H, W, planes.device).reshape(-1)] = False
    keep[:GUARD] = False
    keep[GUARD + ntile_rows:] = False
    return bool((planes[:, keep, :].float() == 0).all())


# ----------------------------------------------------------------------------- split row planes (csrc/gemm_tc.cu)
GEMM_PAD = 256  # rows and K elements of a packed matrix are zero-padded to multiples of this


def pack_rows_split(x: torch.Tensor, rows_cap: int | None = None) -> torch.Tensor:
    """Host-side restatement of ``ddpb_pack_rows_split``: a fp32 matrix [R, K] as the operand format of
    the general tensor-core GEMM — ``[2, Kp/8, Rp, 8]`` bf16 with ``out[0]`` the nearest-bf16 terms,
    ``out[1]`` the bf16-rounded remainders (x = hi + lo), element (r, k) at ``[set, k // 8, r, k % 8]``
    and Rp, Kp the sizes rounded up to ``GEMM_PAD`` (padding zero)."""
    R, K = x.shape
    rp = -(-(rows_cap if rows_cap is not None else R) // GEMM_PAD) * GEMM_PAD
    kp = -(-K // GEMM_PAD) * GEMM_PAD
    hi = x.float().bfloat16()
    lo = (x.float() - hi.float()).bfloat16()
    out = torch.zeros(2, kp // 8, rp, 8, dtype=torch.bfloat16, device=x.device)
    for s, term in enumerate((hi, lo)):
        full = torch.zeros(rp, kp, dtype=torch.bfloat16, device=x.device)
        full[:R, :K] = term
        out[s] = full.view(rp, kp // 8, 8).permute(1, 0, 2)
    return out


def unpack_rows_split(planes: torch.Tensor, R: int, K: int) -> torch.Tensor:
    """[2, Kp/8, Rp, 8] split row planes -> the fp32 matrix [R, K] they represent (hi + lo)."""
    both = planes.float().sum(0)                       # [Kp/8, Rp, 8]
    return both.permute(1, 0, 2).reshape(both.shape[1], -1)[:R, :K].contiguous()
