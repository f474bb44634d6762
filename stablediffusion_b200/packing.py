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


# Nearest-2x upsample folded into the 3x3 conv that follows it (Upsample2D of the UNet up blocks and of the VAE decoder).
# Output pixel (2i+py, 2j+px) of conv3x3(upsample2x(x)) reads x at rows i + floor((py + dh) / 2), dh in {-1, 0, 1}: for py = 0
# kernel row 0 sees x row i-1 and kernel rows 1, 2 both see x row i; for py = 1 rows 0, 1 see x row i and row 2 sees x row
# i+1 (same for columns).  So each of the four output phases is a 2x2-tap conv over the LOW-resolution input whose taps are
# sums of the original kernel entries: 16 tap-GEMMs instead of 36 (2.25x fewer MACs) and no upsampled tensor in HBM.  The
# zero padding of the upsampled image maps exactly onto out-of-bounds rows / columns of the low-resolution input.
_UP_ROWS = {0: ((0,), (1, 2)), 1: ((0, 1), (2,))}     # phase -> kernel rows feeding the two low-resolution shifts
_UP_SHIFT = {0: (-1, 0), 1: (0, 1)}                    # phase -> the two shifts


def upsample_conv_phase_weights(w: torch.Tensor) -> torch.Tensor:
    """[C_out, C_in, 3, 3] -> [4, C_out, 4*C_in]: phase py*2+px, K laid out [tap (row shift, col shift), C_in]; sums in the
    dtype of `w` (fp32 at pack time, rounded to bf16 once afterwards)."""
    co, ci = w.shape[:2]
    out = []
    for py in (0, 1):
        for px in (0, 1):
            taps = [sum(w[:, :, r, s] for r in ra for s in sb) for ra in _UP_ROWS[py] for sb in _UP_ROWS[px]]
            out.append(torch.stack(taps, dim=1).reshape(co, 4 * ci))
    return torch.stack(out).contiguous()


def upsample_conv_phase_taps(py: int, px: int):
    """Tap table (phase, dh, dw) of output phase (py, px), in the K order of upsample_conv_phase_weights."""
    return [(0, dh, dw) for dh in _UP_SHIFT[py] for dw in _UP_SHIFT[px]]
