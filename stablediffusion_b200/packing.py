"""Weight re-layout done once at load time (pure index shuffles; no arithmetic besides the LoRA merge)."""
from __future__ import annotations

import torch


def geglu_block_n(n_total: int) -> int:
    """N tile used for a GEGLU projection of width n_total (= 8*C); must match the library's auto choice
    (csrc/gemm.cu choose_gemm_tiling): block_n/2 output columns = whole 32-column staging blocks of the epilogue."""
    if n_total % 256 == 0:
        return 256
    if n_total % 128 == 0:
        return 128
    raise ValueError(f"GEGLU width {n_total} is not a multiple of 128")


def geglu_interleave(t: torch.Tensor, block_n: int) -> torch.Tensor:
    """Reorder the rows of a GEGLU projection ([h rows ; g rows], diffusers `GEGLU.proj`) so that every
    block_n-row tile holds block_n/2 h rows followed by the matching block_n/2 g rows — the GEMM epilogue
    then computes h*gelu(g) without leaving the tile."""
    n = t.shape[0]
    half = n // 2
    hb = block_n // 2
    if n % block_n != 0:
        raise ValueError(f"{n} rows not divisible by tile {block_n}")
    h = t[:half].reshape(half // hb, hb, *t.shape[1:])
    g = t[half:].reshape(half // hb, hb, *t.shape[1:])
    return torch.cat([h, g], dim=1).reshape(n, *t.shape[1:])


def conv_weight_to_gemm(w: torch.Tensor) -> torch.Tensor:
    """[C_out, C_in, kh, kw] (torch / diffusers) -> [C_out, kh*kw*C_in] with taps outermost, channels contiguous."""
    co, ci, kh, kw = w.shape
    return w.permute(0, 2, 3, 1).reshape(co, kh * kw * ci).contiguous()
