"""Lowering helpers: torch CUDA tensors -> ops of an `_native.Program`.

PyTorch is used here for device memory only (allocation, views, `data_ptr()`); every arithmetic op of the
hot path is one of the library's own kernels.  Tensors handed to the builder must stay alive as long as the
program does — the builder keeps references.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import torch

from . import _native as N

_DT = {torch.bfloat16: N.BF16, torch.float32: N.F32, torch.float16: N.F16}

# tap tables (phase, dh, dw) for the 3x3 convolutions of the path
TAPS_3X3_S1 = [(0, r - 1, s - 1) for r in range(3) for s in range(3)]
# stride 2, padding 1 (UNet Downsample2D): input row 2*oh + r - 1 -> phase row parity / shift
_S2P1 = [(1, -1), (0, 0), (1, 0)]
TAPS_3X3_S2_PAD1 = [(_S2P1[r][0] * 2 + _S2P1[s][0], _S2P1[r][1], _S2P1[s][1]) for r in range(3) for s in range(3)]
# stride 2, padding (0,1,0,1) (VAE encoder Downsample2D): input row 2*oh + r
_S2P0 = [(0, 0), (1, 0), (0, 1)]
TAPS_3X3_S2_PAD0 = [(_S2P0[r][0] * 2 + _S2P0[s][0], _S2P0[r][1], _S2P0[s][1]) for r in range(3) for s in range(3)]
TAPS_1X1 = [(0, 0, 0)]


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def dtype_code(t: torch.Tensor) -> int:
    try:
        return _DT[t.dtype]
    except KeyError:
        raise ValueError(f"unsupported tensor dtype {t.dtype}") from None


def _rows_view(t: torch.Tensor):
    """(data_ptr, rows, cols, row pitch) of a 2-D bf16 view whose last dim is contiguous."""
    if t.dtype != torch.bfloat16:
        raise ValueError(f"expected bf16 activations, got {t.dtype}")
    if t.dim() != 2 or t.stride(1) != 1:
        raise ValueError(f"expected a 2-D row view with contiguous columns, got shape {tuple(t.shape)} strides {t.stride()}")
    return t.data_ptr(), t.shape[0], t.shape[1], t.stride(0)


class Builder:
    """Collects ops into one program; owns scratch workspaces shared by sequential ops."""

    def __init__(self, device: torch.device, workspace_bytes: int = 32 << 20):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("stablediffusion_b200 runs on CUDA (sm_100a) only; there is no CPU path")
        idx = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.device = torch.device("cuda", idx)
        self.prog = N.Program(idx)
        self._ws = torch.empty(workspace_bytes, dtype=torch.uint8, device=self.device)
        self._gn_ws = torch.empty(1 << 20, dtype=torch.uint8, device=self.device)
        self.prog.keep(self._ws, self._gn_ws)
        self.gemm_flops = 0  # algorithmic 2*MAC of the GEMM ops added
        self.attn_flops = 0  # algorithmic 4*B*heads*Nq*Nkv*d of the attention ops added
        self.tag = 0
        self.meta = {}  # op index -> short description (profiling aid)

    @property
    def flops(self):
        return self.gemm_flops + self.attn_flops

    # ---- memory -------------------------------------------------------------------------------------
    def empty(self, shape: Sequence[int], dtype=torch.bfloat16) -> torch.Tensor:
        t = torch.empty(tuple(shape), dtype=dtype, device=self.device)
        self.prog.keep(t)
        return t

    def zeros(self, shape: Sequence[int], dtype=torch.bfloat16) -> torch.Tensor:
        t = torch.zeros(tuple(shape), dtype=dtype, device=self.device)
        self.prog.keep(t)
        return t

    # ---- GEMM family --------------------------------------------------------------------------------
    def _finish_gemm(self, g: N.Gemm, keep) -> None:
        need = N.lib().b200sd_gemm_workspace_bytes(C.byref(g))
        if need < 0:
            N.check(int(need))
        if need > 0:
            ws = self._ws if need <= self._ws.numel() else self.empty([need], torch.uint8)
            g.workspace = ws.data_ptr()
            g.workspace_bytes = ws.numel()
        idx = self.prog.add(N.OP_GEMM, g, self.tag, keep)
        self.meta[idx] = self._pending_meta

    @staticmethod
    def _epilogue(g: N.Gemm, bias, rowvec, residual, ldr, silu, geglu, out_f32, bias_row, alpha, rows_per_group, act=None):
        flags = 0
        if bias is not None:
            if bias.dtype != torch.float32:
                raise ValueError("bias must be fp32")
            flags |= N.GEMM_BIAS_ROW if bias_row else N.GEMM_BIAS
            g.bias = bias.data_ptr()
        if rowvec is not None:
            if rowvec.dtype != torch.float32 or rowvec.dim() != 2 or rowvec.stride(1) != 1:
                raise ValueError("rowvec must be a 2-D fp32 view")
            flags |= N.GEMM_ROWVEC
            g.rowvec = rowvec.data_ptr()
            g.ld_rowvec = rowvec.stride(0)
            g.rows_per_group = rows_per_group
        if residual is not None:
            flags |= N.GEMM_RESIDUAL
            g.residual = residual.data_ptr()
            g.ldr = ldr
        if silu:
            flags |= N.GEMM_SILU
        if geglu:
            flags |= N.GEMM_GEGLU
        if out_f32:
            flags |= N.GEMM_OUT_F32
        if act is not None:
            if act not in ("quick_gelu", "gelu"):
                raise ValueError(f"unknown activation '{act}'")
            flags |= N.GEMM_QUICK_GELU if act == "quick_gelu" else N.GEMM_GELU
        g.flags = flags
        g.alpha = alpha

    def linear(self, a, w, out=None, bias=None, residual=None, rowvec=None, rows_per_group=0, silu=False, geglu=False,
               out_f32=False, bias_row=False, alpha=1.0, a2=None, splits=0, block_n=0, cta_pair=0, act=None, rows_per_sample=0):
        """out[M,N] = a[M,K] @ w[N,K]^T (+ epilogue).  `a2` appends a second K segment ([M,K2], weights w[:, K:]).
        `rows_per_sample`: rows of `a` that belong to one image (only read in the batch-invariant mode)."""
        ap, M, K, lda = _rows_view(a)
        wp, Nn, Kw, ldw = _rows_view(w)
        K2 = 0
        if a2 is not None:
            a2p, M2, K2, lda2 = _rows_view(a2)
            if M2 != M:
                raise ValueError("linear: a2 rows differ")
        if Kw != K + K2:
            raise ValueError(f"linear: weight K {Kw} != activation K {K + K2}")
        n_out = Nn // 2 if geglu else Nn
        if out is None:
            out = self.empty([M, n_out], torch.float32 if out_f32 else torch.bfloat16)
        if out.dim() != 2 or out.shape[0] != M or out.shape[1] != n_out or out.stride(1) != 1:
            raise ValueError(f"linear: out has shape {tuple(out.shape)}, expected {(M, n_out)}")
        g = N.Gemm()
        g.spatial = 0
        g.nseg = 1 if a2 is None else 2
        g.seg[0].a, g.seg[0].w, g.seg[0].lda, g.seg[0].channels, g.seg[0].taps = ap, wp, lda, K, 1
        if a2 is not None:
            g.seg[1].a, g.seg[1].w, g.seg[1].lda, g.seg[1].channels, g.seg[1].taps = a2p, wp + 2 * K, lda2, K2, 1
        g.ldw = ldw
        g.M, g.N = M, Nn
        g.out, g.ldc = out.data_ptr(), out.stride(0)
        g.splits, g.block_n, g.cta_pair = splits, block_n, cta_pair
        g.rows_per_sample = int(rows_per_sample)
        ldr = 0
        if residual is not None:
            _, rm, rn, ldr = _rows_view(residual)
            if rm != M or rn != n_out:
                raise ValueError("linear: residual shape mismatch")
        self._epilogue(g, bias, rowvec, residual, ldr, silu, geglu, out_f32, bias_row, alpha, rows_per_group, act)
        self._pending_meta = dict(op='linear', M=M, N=Nn, K=K + K2, flops=2 * M * Nn * (K + K2), geglu=geglu)
        self._finish_gemm(g, (a, w, out, bias, residual, rowvec, a2))
        self.gemm_flops += 2 * M * Nn * (K + K2)
        return out

    def conv(self, x, w, out=None, taps=TAPS_3X3_S1, phases=1, out_hw=None, bias=None, rowvec=None, residual=None,
             silu=False, extra=None, out_scatter=None, splits=0, block_n=0, cta_pair=0, algo_flops=None):
        """Implicit-GEMM convolution on NHWC bf16.

        x: [NB,H,W,C] (phases == 1) or [4,NB,H/2,W/2,C] (after phase_split); w: [Cout, len(taps)*C (+ extra K)].
        extra: optional list of (x_i [NB,H,W,C_i], 1x1) segments appended along K (their weights follow in w).
        rowvec: fp32 [NB, Cout] added per image (time embedding); residual: [NB,H,W,Cout].
        out_scatter: (OH, OW, sy, sx, oy, ox) to write a strided sub-lattice of a larger output.
        algo_flops: algorithmic 2*MAC this op stands for when that differs from what it executes (folded upsample).
        """
        if x.dtype != torch.bfloat16 or not x.is_contiguous():
            raise ValueError("conv: x must be contiguous bf16")
        if phases == 1:
            NB, Hs, Ws, Cin = x.shape
        else:
            ph, NB, Hs, Ws, Cin = x.shape
            if ph != phases:
                raise ValueError("conv: phase dimension mismatch")
        H, W = out_hw if out_hw is not None else (Hs, Ws)
        wp, Cout, Kw, ldw = _rows_view(w)
        segs = [(x, Cin, list(taps), phases, Hs, Ws)]
        for xe in (extra or []):
            if xe.dtype != torch.bfloat16 or not xe.is_contiguous() or xe.shape[:3] != (NB, H, W):
                raise ValueError("conv: extra segment must be contiguous bf16 [NB,H,W,C]")
            segs.append((xe, xe.shape[3], list(TAPS_1X1), 1, H, W))
        ktot = sum(c * len(t) for _, c, t, _, _, _ in segs)
        if ktot != Kw:
            raise ValueError(f"conv: weight K {Kw} != sum of segments {ktot}")
        if out is None:
            if out_scatter is not None:
                raise ValueError("conv: out_scatter needs the (larger) output tensor")
            out = self.empty([NB, H, W, Cout])
        g = N.Gemm()
        g.spatial = 1
        g.nseg = len(segs)
        koff = 0
        for i, (xs, c, tp, ph, hs, ws) in enumerate(segs):
            s = g.seg[i]
            s.a, s.w, s.lda, s.channels, s.taps = xs.data_ptr(), wp + 2 * koff, c, c, len(tp)
            s.Hs, s.Ws, s.phases = hs, ws, ph
            for j, (p_, dh, dw) in enumerate(tp):
                s.tap_phase[j], s.tap_dh[j], s.tap_dw[j] = p_, dh, dw
            koff += c * len(tp)
        g.ldw = ldw
        g.N = Cout
        g.NB, g.H, g.W = NB, H, W
        g.M = NB * H * W
        g.out = out.data_ptr()
        g.ldc = out.shape[-1]
        if out_scatter is not None:
            g.OH, g.OW, g.sy, g.sx, g.oy, g.ox = out_scatter
        else:
            g.OH, g.OW, g.sy, g.sx, g.oy, g.ox = H, W, 1, 1, 0, 0
        g.splits, g.block_n, g.cta_pair = splits, block_n, cta_pair
        ldr = residual.shape[-1] if residual is not None else 0
        self._epilogue(g, bias, rowvec, residual, ldr, silu, False, False, False, 1.0, 0)
        flops = 2 * NB * H * W * Cout * ktot if algo_flops is None else int(algo_flops)
        self._pending_meta = dict(op=f'conv{len(taps)}', M=NB * H * W, N=Cout, K=ktot, flops=flops, hw=(H, W))
        self._finish_gemm(g, (x, w, out, bias, residual, rowvec) + tuple(extra or ()))
        self.gemm_flops += flops
        return out

    def upsample2x_conv(self, x, w, w_phase, bias=None, fold=None):
        """conv3x3(nearest_upsample_2x(x)) (+bias): x [NB,H,W,C] -> [NB,2H,2W,Cout].
        w: [Cout, 9C] plain packing; w_phase: [4, Cout, 4C] (packing.upsample_conv_phase_weights).
        Folded form (four 2x2-tap phase convs over x, each writing its lattice of the output: 2.25x fewer MACs, no upsampled
        tensor) where the phase GEMMs are large enough to fill the GPU, else upsample2x + one 3x3 conv (fewer launches).
        `fold`: None = decide here; the decision reads the batch only outside the batch-invariant mode (the two forms round
        differently: phase weights are pre-summed)."""
        from .packing import upsample_conv_phase_taps

        NB, H, W, Cin = x.shape
        Cout = w.shape[0]
        if fold is None:
            rows = (8 if N.batch_invariant() else NB) * H * W      # kCanonBatch images in the batch-invariant mode
            fold = rows >= 4096
        if not fold:
            return self.conv(self.upsample2x(x), w, bias=bias)
        out = self.empty([NB, 2 * H, 2 * W, Cout])
        for py in (0, 1):
            for px in (0, 1):
                self.conv(x, w_phase[py * 2 + px], out=out, taps=upsample_conv_phase_taps(py, px), bias=bias,
                          out_scatter=(2 * H, 2 * W, 2, 2, py, px), algo_flops=2 * NB * H * W * Cout * 9 * Cin)
        return out

    # ---- normalisation ------------------------------------------------------------------------------
    def groupnorm(self, x1, gamma, beta, eps, silu, x2=None, groups=32, out=None):
        """x1 [NB,HW,C1] (+ x2 [NB,HW,C2] concatenated along channels) -> bf16 [NB,HW,C1+C2]."""
        NB, HW, C1 = x1.shape
        C2 = 0 if x2 is None else x2.shape[2]
        if out is None:
            out = self.empty([NB, HW, C1 + C2])
        chunks = N.lib().b200sd_groupnorm_chunks(NB, HW, C1 + C2)
        need = NB * chunks * groups * 2 * 4
        if need > self._gn_ws.numel():  # (sequential ops share one partials buffer; it only ever grows)
            self._gn_ws = torch.empty(need, dtype=torch.uint8, device=self.device)
            self.prog.keep(self._gn_ws)
        p = N.GroupNorm()
        p.x1, p.ld1, p.C1 = x1.data_ptr(), x1.stride(1), C1
        if x2 is not None:
            p.x2, p.ld2, p.C2 = x2.data_ptr(), x2.stride(1), C2
        p.NB, p.HW, p.groups, p.silu, p.eps = NB, HW, groups, int(bool(silu)), float(eps)
        p.gamma, p.beta, p.out = gamma.data_ptr(), beta.data_ptr(), out.data_ptr()
        p.partials, p.partials_bytes = self._gn_ws.data_ptr(), self._gn_ws.numel()
        idx = self.prog.add(N.OP_GROUPNORM, p, self.tag, (x1, x2, gamma, beta, out))
        self.meta[idx] = dict(op='groupnorm', NB=NB, HW=HW, C=C1 + C2, bytes=4 * NB * HW * (C1 + C2))
        return out

    def layernorm(self, x, gamma, beta, eps, out=None):
        xp, rows, Cc, ldx = _rows_view(x)
        if out is None:
            out = self.empty([rows, Cc])
        p = N.LayerNorm()
        p.x, p.out, p.ldx, p.ldo, p.rows, p.C, p.eps = xp, out.data_ptr(), ldx, out.stride(0), rows, Cc, float(eps)
        p.gamma, p.beta = gamma.data_ptr(), beta.data_ptr()
        idx = self.prog.add(N.OP_LAYERNORM, p, self.tag, (x, gamma, beta, out))
        self.meta[idx] = dict(op='layernorm', rows=rows, C=Cc, bytes=4 * rows * Cc)
        return out

    # ---- attention ----------------------------------------------------------------------------------
    def attention(self, q, k, v, heads, d, scale, out=None, causal=False):
        """q [B,Nq,*], k/v [B,Nkv,*]: 3-D views, last dim contiguous, batch stride = rows*pitch."""
        B, Nq, _ = q.shape
        Nkv = k.shape[1]
        for t, rows in ((q, Nq), (k, Nkv), (v, Nkv)):
            if t.dtype != torch.bfloat16 or t.stride(2) != 1 or (B > 1 and t.stride(0) != rows * t.stride(1)):
                raise ValueError("attention: operands must be bf16 [B,rows,cols] views with dense batch stride")
        if out is None:
            out = self.empty([B, Nq, heads * d])
        p = N.Attention()
        p.q, p.k, p.v, p.out = q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr()
        p.ldq, p.ldk, p.ldv, p.ldo = q.stride(1), k.stride(1), v.stride(1), out.stride(1)
        p.B, p.heads, p.Nq, p.Nkv, p.d, p.scale = B, heads, Nq, Nkv, d, float(scale)
        p.causal = 1 if causal else 0
        idx = self.prog.add(N.OP_ATTENTION, p, self.tag, (q, k, v, out))
        self.meta[idx] = dict(op='attention', B=B, heads=heads, Nq=Nq, Nkv=Nkv, d=d, flops=4 * B * heads * Nq * Nkv * d)
        self.attn_flops += 4 * B * heads * Nq * Nkv * d
        return out

    def softmax_rows(self, x, scale, out=None):
        xp, rows, cols, ldx = _rows_view(x)
        if out is None:
            out = self.empty([rows, cols])
        p = N.SoftmaxRows()
        p.x, p.out, p.ldx, p.ldo, p.rows, p.cols, p.scale = xp, out.data_ptr(), ldx, out.stride(0), rows, cols, float(scale)
        self.prog.add(N.OP_SOFTMAX_ROWS, p, self.tag, (x, out))
        return out

    # ---- small ops ----------------------------------------------------------------------------------
    def timestep_embed(self, rows, dim, out=None, t=None, t_stride=0, step_table=None, step_counter=None,
                       flip_sin_to_cos=True, freq_shift=0.0):
        if out is None:
            out = self.empty([rows, dim])
        p = N.TimestepEmbed()
        p.t = ptr(t)
        p.step_table, p.step_counter = ptr(step_table), ptr(step_counter)
        p.t_stride, p.rows, p.dim, p.flip_sin_to_cos = t_stride, rows, dim, int(flip_sin_to_cos)
        p.freq_shift, p.max_period, p.out = float(freq_shift), 10000.0, out.data_ptr()
        self.prog.add(N.OP_TIMESTEP_EMBED, p, self.tag, (t, step_table, step_counter, out))
        return out

    def prep_latents(self, latents, out, rep, fixed_scale=1.0, step_table=None, step_counter=None):
        B, Cc, h, w = latents.shape
        if not latents.is_contiguous():
            raise ValueError("prep_latents: latents must be contiguous NCHW")
        p = N.PrepLatents()
        p.latents, p.out = latents.data_ptr(), out.data_ptr()
        p.step_table, p.step_counter = ptr(step_table), ptr(step_counter)
        p.fixed_scale, p.in_dtype = float(fixed_scale), dtype_code(latents)
        p.B, p.C, p.hw, p.Cpad, p.rep = B, Cc, h * w, out.shape[-1], rep
        self.prog.add(N.OP_PREP_LATENTS, p, self.tag, (latents, out, step_table, step_counter))
        return out

    def upsample2x(self, x, out=None):
        NB, H, W, Cc = x.shape
        if out is None:
            out = self.empty([NB, 2 * H, 2 * W, Cc])
        p = N.Upsample2x()
        p.x, p.out, p.NB, p.H, p.W, p.C = x.data_ptr(), out.data_ptr(), NB, H, W, Cc
        self.prog.add(N.OP_UPSAMPLE2X, p, self.tag, (x, out))
        return out

    def phase_split(self, x, out=None):
        NB, H, W, Cc = x.shape
        if out is None:
            out = self.empty([4, NB, H // 2, W // 2, Cc])
        p = N.PhaseSplit()
        p.x, p.out, p.NB, p.H, p.W, p.C = x.data_ptr(), out.data_ptr(), NB, H, W, Cc
        self.prog.add(N.OP_PHASE_SPLIT, p, self.tag, (x, out))
        return out

    def cfg_step(self, eps, x, x_out, step_table, step_counter, guidance_scale, cfg, nsteps, x0_prev=None, next_in=None, rep=2,
                 gs_from_table=False, noise=None, eps_hist=None, x_saved=None, blend=None):
        B, Cc, h, w = x.shape
        p = N.CfgStep()
        p.eps, p.ld_eps = eps.data_ptr(), eps.shape[-1]
        p.x, p.x_out, p.x0_prev, p.next_in = x.data_ptr(), x_out.data_ptr(), ptr(x0_prev), ptr(next_in)
        p.step_table, p.step_counter = step_table.data_ptr(), step_counter.data_ptr()
        p.guidance_scale, p.cfg, p.x_dtype = float(guidance_scale), int(bool(cfg)), dtype_code(x)
        p.B, p.C, p.hw = B, Cc, h * w
        p.Cpad = next_in.shape[-1] if next_in is not None else 0
        p.rep, p.nsteps, p.gs_from_table = rep, nsteps, int(bool(gs_from_table))
        for t in (noise, eps_hist, x_saved):
            if t is not None and (t.dtype != torch.float32 or not t.is_contiguous()):
                raise ValueError("cfg_step: noise / eps_hist / x_saved must be contiguous fp32")
        p.noise, p.eps_hist, p.x_saved = ptr(noise), ptr(eps_hist), ptr(x_saved)
        if blend is not None:  # (mask [B, hw], image latents [B, C, hw], start noise [B, C, hw]), all fp32
            for t in blend:
                if t.dtype != torch.float32 or not t.is_contiguous():
                    raise ValueError("cfg_step: blend tensors must be contiguous fp32")
            p.blend_mask, p.blend_image, p.blend_noise = (t.data_ptr() for t in blend)
            self.prog.keep(*blend)
        self.prog.add(N.OP_CFG_STEP, p, self.tag, (eps, x, x_out, step_table, step_counter, x0_prev, next_in, noise, eps_hist, x_saved))
        return x_out

    def axpby(self, x, a, y=None, b=0.0, out=None, coef=None):
        """out = a*x + b*y.  `coef`: optional fp32 device tensor {a, b} read at run time (overrides the baked a, b)."""
        if out is None:
            out = self.empty(x.shape, x.dtype)
        p = N.Axpby()
        p.coef = ptr(coef)
        p.x, p.y, p.out, p.a, p.b = x.data_ptr(), ptr(y), out.data_ptr(), float(a), float(b)
        p.x_dtype, p.y_dtype, p.out_dtype = dtype_code(x), (dtype_code(y) if y is not None else 0), dtype_code(out)
        p.n = x.numel()
        self.prog.add(N.OP_AXPBY, p, self.tag, (x, y, out, coef))
        return out

    def image_out(self, x, C_valid, out):
        B, HW, ldx = x.shape
        p = N.ImageOut()
        p.x, p.out, p.ldx, p.B, p.C, p.HW, p.out_dtype = x.data_ptr(), out.data_ptr(), ldx, B, C_valid, HW, dtype_code(out)
        self.prog.add(N.OP_IMAGE_OUT, p, self.tag, (x, out))
        return out

    def image_u8(self, x, out):
        """x: [B, C, H*W] image in [-1, 1] (bf16 / fp16 / fp32) -> out: uint8 [B, H*W, C] (handler_logic.py:21-29)."""
        B, Cc, HW = x.shape
        p = N.ImageU8()
        p.x, p.out, p.B, p.C, p.HW, p.x_dtype = x.data_ptr(), out.data_ptr(), B, Cc, HW, dtype_code(x)
        self.prog.add(N.OP_IMAGE_U8, p, self.tag, (x, out))
        return out

    def embed(self, ids, tok, pos, out=None):
        """ids: int32 [rows]; tok: bf16 [vocab, D]; pos: bf16 [L, D] -> bf16 [rows, D] = tok[ids] + pos[row % L]."""
        if ids.dtype != torch.int32 or tok.dtype != torch.bfloat16 or pos.dtype != torch.bfloat16:
            raise ValueError("embed: ids must be int32, tables bf16")
        rows, (vocab, D), Lp = ids.numel(), tok.shape, pos.shape[0]
        if out is None:
            out = self.empty([rows, D])
        p = N.Embed()
        p.ids, p.tok, p.pos, p.out = ids.data_ptr(), tok.data_ptr(), pos.data_ptr(), out.data_ptr()
        p.rows, p.L, p.D, p.vocab = rows, Lp, D, vocab
        self.prog.add(N.OP_EMBED, p, self.tag, (ids, tok, pos, out))
        return out

    def gauss_sample(self, moments, noise, out, scale=1.0):
        """moments [B, 2C, h, w] (mean | logvar), noise [B, C, h, w] -> out [B, C, h, w] = scale*(mean + std*noise)."""
        B, C2, h, w = moments.shape
        if C2 % 2 or noise.shape != (B, C2 // 2, h, w) or out.shape != noise.shape:
            raise ValueError("gauss_sample: shapes must be [B,2C,h,w], [B,C,h,w], [B,C,h,w]")
        for t in (moments, noise, out):
            if not t.is_contiguous():
                raise ValueError("gauss_sample: tensors must be contiguous NCHW")
        p = N.GaussSample()
        p.moments, p.noise, p.out, p.scale = moments.data_ptr(), noise.data_ptr(), out.data_ptr(), float(scale)
        p.m_dtype, p.n_dtype, p.out_dtype = dtype_code(moments), dtype_code(noise), dtype_code(out)
        p.B, p.C, p.hw = B, C2 // 2, h * w
        self.prog.add(N.OP_GAUSS_SAMPLE, p, self.tag, (moments, noise, out))
        return out

    def step_advance(self, step_counter, ack_remote: int = 0, state=None):
        """++step counter; with the pair split also the acknowledgement to the peer (`ack_remote`: raw peer address)."""
        p = N.StepAdvance()
        p.step_counter = step_counter.data_ptr()
        if ack_remote:
            p.ack_remote, p.state = int(ack_remote), state.data_ptr()
        self.prog.add(N.OP_STEP_ADVANCE, p, self.tag, (step_counter, state))

    def pair_exchange(self, src, dst_remote: int, sync_local: int, sync_remote: int, state, timeout_ms: int = 20000):
        """This rank's eps half (`src`, a tensor inside the local window) -> the same slot of the peer's window
        (`dst_remote`, raw peer address), flag / ack handshake through the windows' sync words (csrc/peer.cu)."""
        if not src.is_contiguous():
            raise ValueError("pair_exchange: src must be contiguous")
        p = N.PairExchange()
        p.src, p.dst_remote, p.bytes = src.data_ptr(), int(dst_remote), src.numel() * src.element_size()
        p.sync_local, p.sync_remote, p.state, p.timeout_ms = int(sync_local), int(sync_remote), state.data_ptr(), int(timeout_ms)
        self.prog.add(N.OP_PAIR_EXCHANGE, p, self.tag, (src, state))
