"""B200 UNet: the replacement for `model.base` (diffusers `UNet2DConditionModel`) on the reference's hot
path — call site pipelines/sd_unified_pipeline.py:475-482, loader models/stable_diffusion.py:117-123.

The network is *lowered once* to a flat op program (stablediffusion_b200/lowering.py -> libb200sd.so):
weights are re-laid-out for the implicit-GEMM kernel at load (LoRA already merged), activations are NHWC
bf16 end to end (token-major == NHWC, so resnet -> transformer -> resnet needs no permutes), and the fusions
listed in SURVEY §2.3 are applied structurally:
  * q/k/v of self-attention are one GEMM (N = 3C); every cross-attention k/v projection of the network is
    ONE GEMM over the prompt context (N = sum 2C); every resnet's time_emb_proj is ONE GEMM (N = sum C_out)
  * GroupNorm+SiLU is one pass that also performs the skip-connection concat (two-source read)
  * conv1 adds bias + time-embedding row vector in its epilogue; conv2 adds bias + residual, and when the
    block has a 1x1 `conv_shortcut` that shortcut is folded into conv2's K loop as extra segments
  * GEGLU (h * gelu(g)) is the epilogue of the 8C projection; every residual add is a GEMM epilogue
State-dict key names are diffusers' (scripts/convert_from_A1111.py:240-485), so real checkpoints load.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Dict, List, Optional, Tuple

import torch

from . import _native as N
from . import lowering as L
from .config import Config, unet_config
from .packing import conv_weight_to_gemm, geglu_block_n, geglu_interleave, upsample_conv_phase_weights

TEMB_MULT = 4


def _tup(v, n):
    return tuple(v) if isinstance(v, (tuple, list)) else (v,) * n


# ------------------------------------------------------------------------------------------------------
# structure: one place that knows the topology; parameter shapes and the op program both derive from it
# ------------------------------------------------------------------------------------------------------
def unet_structure(cfg: Config) -> dict:
    boc = tuple(cfg["block_out_channels"])
    nb = len(boc)
    heads = _tup(cfg.get("num_attention_heads") or cfg["attention_head_dim"], nb)  # "attention_head_dim" == heads
    tl = _tup(cfg["transformer_layers_per_block"], nb)
    lpb = cfg["layers_per_block"]
    s = dict(down=[], up=[], temb=boc[0] * TEMB_MULT)
    out_ch = boc[0]
    for i, t in enumerate(cfg["down_block_types"]):
        in_ch, out_ch = out_ch, boc[i]
        blk = dict(name=f"down_blocks.{i}", resnets=[], attn=None, down=(i != nb - 1), ch=out_ch)
        for j in range(lpb):
            blk["resnets"].append((in_ch if j == 0 else out_ch, 0, out_ch))
        if t.startswith("CrossAttn"):
            blk["attn"] = dict(heads=heads[i], layers=tl[i])
        s["down"].append(blk)
    s["mid"] = dict(name="mid_block", ch=boc[-1], attn=dict(heads=heads[-1], layers=tl[-1]))
    rev = list(reversed(boc))
    out_ch = rev[0]
    for i, t in enumerate(cfg["up_block_types"]):
        prev, out_ch = out_ch, rev[i]
        in_ch = rev[min(i + 1, nb - 1)]
        blk = dict(name=f"up_blocks.{i}", resnets=[], attn=None, up=(i != nb - 1), ch=out_ch)
        for j in range(lpb + 1):
            skip = in_ch if j == lpb else out_ch
            rin = prev if j == 0 else out_ch
            blk["resnets"].append((rin, skip, out_ch))
        if t.startswith("CrossAttn"):
            blk["attn"] = dict(heads=heads[nb - 1 - i], layers=tl[nb - 1 - i])
        s["up"].append(blk)
    return s


def _resnet_shapes(p: str, cin: int, cout: int, temb: Optional[int], sh: dict):
    sh[f"{p}.norm1.weight"] = (cin,); sh[f"{p}.norm1.bias"] = (cin,)
    sh[f"{p}.conv1.weight"] = (cout, cin, 3, 3); sh[f"{p}.conv1.bias"] = (cout,)
    if temb is not None:
        sh[f"{p}.time_emb_proj.weight"] = (cout, temb); sh[f"{p}.time_emb_proj.bias"] = (cout,)
    sh[f"{p}.norm2.weight"] = (cout,); sh[f"{p}.norm2.bias"] = (cout,)
    sh[f"{p}.conv2.weight"] = (cout, cout, 3, 3); sh[f"{p}.conv2.bias"] = (cout,)
    if cin != cout:
        sh[f"{p}.conv_shortcut.weight"] = (cout, cin, 1, 1); sh[f"{p}.conv_shortcut.bias"] = (cout,)


def _transformer_shapes(p: str, C: int, layers: int, ctx: int, linear_proj: bool, sh: dict):
    sh[f"{p}.norm.weight"] = (C,); sh[f"{p}.norm.bias"] = (C,)
    pw = (C, C) if linear_proj else (C, C, 1, 1)
    sh[f"{p}.proj_in.weight"] = pw; sh[f"{p}.proj_in.bias"] = (C,)
    for k in range(layers):
        b = f"{p}.transformer_blocks.{k}"
        for n in ("norm1", "norm2", "norm3"):
            sh[f"{b}.{n}.weight"] = (C,); sh[f"{b}.{n}.bias"] = (C,)
        for a, kd in (("attn1", C), ("attn2", ctx)):
            sh[f"{b}.{a}.to_q.weight"] = (C, C)
            sh[f"{b}.{a}.to_k.weight"] = (C, kd)
            sh[f"{b}.{a}.to_v.weight"] = (C, kd)
            sh[f"{b}.{a}.to_out.0.weight"] = (C, C); sh[f"{b}.{a}.to_out.0.bias"] = (C,)
        sh[f"{b}.ff.net.0.proj.weight"] = (8 * C, C); sh[f"{b}.ff.net.0.proj.bias"] = (8 * C,)
        sh[f"{b}.ff.net.2.weight"] = (C, 4 * C); sh[f"{b}.ff.net.2.bias"] = (C,)
    sh[f"{p}.proj_out.weight"] = pw; sh[f"{p}.proj_out.bias"] = (C,)


def unet_param_shapes(cfg: Config) -> "OrderedDict[str, Tuple[int, ...]]":
    s = unet_structure(cfg)
    boc = cfg["block_out_channels"]
    temb, ctx, lin = s["temb"], cfg["cross_attention_dim"], cfg["use_linear_projection"]
    sh: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict()
    sh["conv_in.weight"] = (boc[0], cfg["in_channels"], 3, 3); sh["conv_in.bias"] = (boc[0],)
    sh["time_embedding.linear_1.weight"] = (temb, boc[0]); sh["time_embedding.linear_1.bias"] = (temb,)
    sh["time_embedding.linear_2.weight"] = (temb, temb); sh["time_embedding.linear_2.bias"] = (temb,)
    if cfg.get("addition_embed_type") == "text_time":
        pin = cfg["projection_class_embeddings_input_dim"]
        sh["add_embedding.linear_1.weight"] = (temb, pin); sh["add_embedding.linear_1.bias"] = (temb,)
        sh["add_embedding.linear_2.weight"] = (temb, temb); sh["add_embedding.linear_2.bias"] = (temb,)
    for blk in s["down"]:
        for j, (cin, _, cout) in enumerate(blk["resnets"]):
            _resnet_shapes(f"{blk['name']}.resnets.{j}", cin, cout, temb, sh)
            if blk["attn"]:
                _transformer_shapes(f"{blk['name']}.attentions.{j}", cout, blk["attn"]["layers"], ctx, lin, sh)
        if blk["down"]:
            sh[f"{blk['name']}.downsamplers.0.conv.weight"] = (blk["ch"], blk["ch"], 3, 3)
            sh[f"{blk['name']}.downsamplers.0.conv.bias"] = (blk["ch"],)
    m = s["mid"]
    _resnet_shapes("mid_block.resnets.0", m["ch"], m["ch"], temb, sh)
    _transformer_shapes("mid_block.attentions.0", m["ch"], m["attn"]["layers"], ctx, lin, sh)
    _resnet_shapes("mid_block.resnets.1", m["ch"], m["ch"], temb, sh)
    for blk in s["up"]:
        for j, (rin, skip, cout) in enumerate(blk["resnets"]):
            _resnet_shapes(f"{blk['name']}.resnets.{j}", rin + skip, cout, temb, sh)
            if blk["attn"]:
                _transformer_shapes(f"{blk['name']}.attentions.{j}", cout, blk["attn"]["layers"], ctx, lin, sh)
        if blk["up"]:
            sh[f"{blk['name']}.upsamplers.0.conv.weight"] = (blk["ch"], blk["ch"], 3, 3)
            sh[f"{blk['name']}.upsamplers.0.conv.bias"] = (blk["ch"],)
    sh["conv_norm_out.weight"] = (boc[0],); sh["conv_norm_out.bias"] = (boc[0],)
    sh["conv_out.weight"] = (cfg["out_channels"], boc[0], 3, 3); sh["conv_out.bias"] = (cfg["out_channels"],)
    return sh


# ------------------------------------------------------------------------------------------------------
# weight packing (state dict -> device tensors in GEMM layout)
# ------------------------------------------------------------------------------------------------------
class PackedUNet:
    """Device-resident weights in the layout the kernels consume.  Built from a (LoRA-merged) state dict."""

    def __init__(self, cfg: Config, sd: Dict[str, torch.Tensor], device: torch.device):
        self.cfg = cfg
        self.device = device
        self.s = unet_structure(cfg)
        shapes = unet_param_shapes(cfg)
        missing = [k for k in shapes if k not in sd]
        if missing:
            raise ValueError(f"UNet state dict is missing {len(missing)} tensors, e.g. {missing[:3]}")
        for k, shp in shapes.items():
            got = tuple(sd[k].shape)
            if got != tuple(shp) and not (len(shp) == 4 and shp[2:] == (1, 1) and got == shp[:2]) \
                    and not (len(shp) == 2 and got == tuple(shp) + (1, 1)):
                raise ValueError(f"UNet tensor {k} has shape {got}, expected {tuple(shp)}")
        self.t: Dict[str, torch.Tensor] = {}
        self._sd = sd
        self._pack()
        self._sd = None

    # helpers ---------------------------------------------------------------------------------------
    def _w(self, name):  # master weight as fp32 on the packing device
        return self._sd[name].to(device=self.device, dtype=torch.float32)

    def _bf(self, t):
        return t.to(torch.bfloat16).contiguous()

    def _f32(self, t):
        return t.to(torch.float32).contiguous()

    def _mat(self, name):  # Linear weight or 1x1 conv weight -> [N, K]
        w = self._w(name)
        return w.reshape(w.shape[0], w.shape[1]) if w.dim() == 4 else w

    def _pack_resnet(self, p: str, has_shortcut: bool, temb_rows: list, temb_bias: list):
        T = self.t
        T[f"{p}.norm1.g"], T[f"{p}.norm1.b"] = self._f32(self._w(f"{p}.norm1.weight")), self._f32(self._w(f"{p}.norm1.bias"))
        T[f"{p}.norm2.g"], T[f"{p}.norm2.b"] = self._f32(self._w(f"{p}.norm2.weight")), self._f32(self._w(f"{p}.norm2.bias"))
        T[f"{p}.conv1.w"] = self._bf(conv_weight_to_gemm(self._w(f"{p}.conv1.weight")))
        T[f"{p}.conv1.b"] = self._f32(self._w(f"{p}.conv1.bias"))
        w2 = conv_weight_to_gemm(self._w(f"{p}.conv2.weight"))
        b2 = self._w(f"{p}.conv2.bias")
        if has_shortcut:
            w2 = torch.cat([w2, self._mat(f"{p}.conv_shortcut.weight")], dim=1)
            b2 = b2 + self._w(f"{p}.conv_shortcut.bias")
        T[f"{p}.conv2.w"], T[f"{p}.conv2.b"] = self._bf(w2), self._f32(b2)
        off = sum(r.shape[0] for r in temb_rows)
        temb_rows.append(self._w(f"{p}.time_emb_proj.weight"))
        temb_bias.append(self._w(f"{p}.time_emb_proj.bias"))
        return off

    def _pack_transformer(self, p: str, C: int, layers: int, kv_rows: list):
        T = self.t
        T[f"{p}.norm.g"], T[f"{p}.norm.b"] = self._f32(self._w(f"{p}.norm.weight")), self._f32(self._w(f"{p}.norm.bias"))
        for nm in ("proj_in", "proj_out"):
            T[f"{p}.{nm}.w"], T[f"{p}.{nm}.b"] = self._bf(self._mat(f"{p}.{nm}.weight")), self._f32(self._w(f"{p}.{nm}.bias"))
        kv_offs = []
        for k in range(layers):
            b = f"{p}.transformer_blocks.{k}"
            for n in ("norm1", "norm2", "norm3"):
                T[f"{b}.{n}.g"], T[f"{b}.{n}.b"] = self._f32(self._w(f"{b}.{n}.weight")), self._f32(self._w(f"{b}.{n}.bias"))
            T[f"{b}.attn1.qkv.w"] = self._bf(torch.cat([self._w(f"{b}.attn1.to_q.weight"), self._w(f"{b}.attn1.to_k.weight"),
                                                         self._w(f"{b}.attn1.to_v.weight")], dim=0))
            T[f"{b}.attn2.q.w"] = self._bf(self._w(f"{b}.attn2.to_q.weight"))
            off = sum(r.shape[0] for r in kv_rows)
            kv_rows.append(self._w(f"{b}.attn2.to_k.weight"))
            kv_rows.append(self._w(f"{b}.attn2.to_v.weight"))
            kv_offs.append(off)
            for a in ("attn1", "attn2"):
                T[f"{b}.{a}.out.w"] = self._bf(self._w(f"{b}.{a}.to_out.0.weight"))
                T[f"{b}.{a}.out.b"] = self._f32(self._w(f"{b}.{a}.to_out.0.bias"))
            bn = geglu_block_n(8 * C)
            T[f"{b}.ff1.w"] = self._bf(geglu_interleave(self._w(f"{b}.ff.net.0.proj.weight"), bn))
            T[f"{b}.ff1.b"] = self._f32(geglu_interleave(self._w(f"{b}.ff.net.0.proj.bias"), bn))
            T[f"{b}.ff2.w"] = self._bf(self._w(f"{b}.ff.net.2.weight"))
            T[f"{b}.ff2.b"] = self._f32(self._w(f"{b}.ff.net.2.bias"))
        return kv_offs

    def _pack(self):
        cfg, s, T = self.cfg, self.s, self.t
        temb_rows, temb_bias, kv_rows = [], [], []
        self.temb_off: Dict[str, int] = {}
        self.kv_off: Dict[str, List[int]] = {}
        # conv_in: pad C_in (4) to 8 so that a pixel row is 16 bytes (TMA stride granularity)
        w = self._w("conv_in.weight")
        co, ci = w.shape[0], w.shape[1]
        self.cin_pad = (ci + 7) // 8 * 8
        wp = torch.zeros(co, 3, 3, self.cin_pad, device=self.device)
        wp[..., :ci] = w.permute(0, 2, 3, 1)
        T["conv_in.w"], T["conv_in.b"] = self._bf(wp.reshape(co, 9 * self.cin_pad)), self._f32(self._w("conv_in.bias"))
        for nm in ("time_embedding.linear_1", "time_embedding.linear_2"):
            T[f"{nm}.w"], T[f"{nm}.b"] = self._bf(self._w(f"{nm}.weight")), self._f32(self._w(f"{nm}.bias"))
        if cfg.get("addition_embed_type") == "text_time":
            for nm in ("add_embedding.linear_1", "add_embedding.linear_2"):
                T[f"{nm}.w"], T[f"{nm}.b"] = self._bf(self._w(f"{nm}.weight")), self._f32(self._w(f"{nm}.bias"))

        def do_block(blk, resnets):
            for j, (cin, cout) in enumerate(resnets):
                p = f"{blk['name']}.resnets.{j}"
                self.temb_off[p] = self._pack_resnet(p, cin != cout, temb_rows, temb_bias)
                if blk.get("attn"):
                    pa = f"{blk['name']}.attentions.{j}"
                    self.kv_off[pa] = self._pack_transformer(pa, cout, blk["attn"]["layers"], kv_rows)

        for blk in s["down"]:
            do_block(blk, [(cin, cout) for cin, _, cout in blk["resnets"]])
            if blk["down"]:
                p = f"{blk['name']}.downsamplers.0.conv"
                T[f"{p}.w"], T[f"{p}.b"] = self._bf(conv_weight_to_gemm(self._w(f"{p}.weight"))), self._f32(self._w(f"{p}.bias"))
        m = s["mid"]
        self.temb_off["mid_block.resnets.0"] = self._pack_resnet("mid_block.resnets.0", False, temb_rows, temb_bias)
        self.kv_off["mid_block.attentions.0"] = self._pack_transformer("mid_block.attentions.0", m["ch"], m["attn"]["layers"], kv_rows)
        self.temb_off["mid_block.resnets.1"] = self._pack_resnet("mid_block.resnets.1", False, temb_rows, temb_bias)
        for blk in s["up"]:
            do_block(blk, [(rin + skip, cout) for rin, skip, cout in blk["resnets"]])
            if blk["up"]:
                p = f"{blk['name']}.upsamplers.0.conv"
                T[f"{p}.w"], T[f"{p}.b"] = self._bf(conv_weight_to_gemm(self._w(f"{p}.weight"))), self._f32(self._w(f"{p}.bias"))
                T[f"{p}.wph"] = self._bf(upsample_conv_phase_weights(self._w(f"{p}.weight")))
        T["conv_norm_out.g"], T["conv_norm_out.b"] = self._f32(self._w("conv_norm_out.weight")), self._f32(self._w("conv_norm_out.bias"))
        T["conv_out.w"] = self._bf(conv_weight_to_gemm(self._w("conv_out.weight")))
        T["conv_out.b"] = self._f32(self._w("conv_out.bias"))
        T["temb_all.w"], T["temb_all.b"] = self._bf(torch.cat(temb_rows, 0)), self._f32(torch.cat(temb_bias, 0))
        T["ctx_kv.w"] = self._bf(torch.cat(kv_rows, 0))
        self.nbytes = sum(t.numel() * t.element_size() for t in T.values())


# ------------------------------------------------------------------------------------------------------
# lowering the forward pass to ops
# ------------------------------------------------------------------------------------------------------
class UNetLowering:
    """Emits the ops of one UNet forward for a fixed (NB, h, w) into a Builder."""

    def __init__(self, b: L.Builder, pk: PackedUNet, NB: int, h: int, w: int, ctx_len: int):
        self.b, self.pk, self.NB, self.h, self.w, self.ctx_len = b, pk, NB, h, w, ctx_len
        self.cfg, self.s, self.T = pk.cfg, pk.s, pk.t

    # one resnet: x [NB,H,W,C1] (+ skip [NB,H,W,C2]) -> [NB,H,W,Cout]
    def resnet(self, p, x, skip, cout, eps):
        b, T = self.b, self.T
        NB, H, W, C1 = x.shape
        C2 = 0 if skip is None else skip.shape[3]
        h1 = b.groupnorm(x.view(NB, H * W, C1), T[f"{p}.norm1.g"], T[f"{p}.norm1.b"], eps, True,
                         x2=None if skip is None else skip.view(NB, H * W, C2))
        off = self.pk.temb_off[p]
        h2 = b.conv(h1.view(NB, H, W, C1 + C2), T[f"{p}.conv1.w"], bias=T[f"{p}.conv1.b"],
                    rowvec=self.temb_all[:, off:off + cout])
        h3 = b.groupnorm(h2.view(NB, H * W, cout), T[f"{p}.norm2.g"], T[f"{p}.norm2.b"], eps, True)
        h3 = h3.view(NB, H, W, cout)
        if C1 + C2 != cout:
            extra = [x] if skip is None else [x, skip]
            return b.conv(h3, T[f"{p}.conv2.w"], bias=T[f"{p}.conv2.b"], extra=extra)
        return b.conv(h3, T[f"{p}.conv2.w"], bias=T[f"{p}.conv2.b"], residual=x)

    def transformer(self, p, x, heads, layers):
        b, T = self.b, self.T
        NB, H, W, C = x.shape
        tok = NB * H * W
        d = C // heads
        hn = b.groupnorm(x.view(NB, H * W, C), T[f"{p}.norm.g"], T[f"{p}.norm.b"], 1e-6, False)
        rps = H * W  # rows per image (batch-invariant tiling hint)
        t = b.linear(hn.view(tok, C), T[f"{p}.proj_in.w"], bias=T[f"{p}.proj_in.b"], rows_per_sample=rps)
        for k in range(layers):
            blk = f"{p}.transformer_blocks.{k}"
            n1 = b.layernorm(t, T[f"{blk}.norm1.g"], T[f"{blk}.norm1.b"], 1e-5)
            qkv = b.linear(n1, T[f"{blk}.attn1.qkv.w"], rows_per_sample=rps).view(NB, H * W, 3 * C)
            a1 = b.attention(qkv[:, :, :C], qkv[:, :, C:2 * C], qkv[:, :, 2 * C:], heads, d, d ** -0.5)
            b.linear(a1.view(tok, C), T[f"{blk}.attn1.out.w"], bias=T[f"{blk}.attn1.out.b"], residual=t, out=t, rows_per_sample=rps)
            n2 = b.layernorm(t, T[f"{blk}.norm2.g"], T[f"{blk}.norm2.b"], 1e-5)
            q2 = b.linear(n2, T[f"{blk}.attn2.q.w"], rows_per_sample=rps).view(NB, H * W, C)
            off = self.pk.kv_off[p][k]
            kv = self.ctx_kv.view(NB, self.ctx_len, -1)
            a2 = b.attention(q2, kv[:, :, off:off + C], kv[:, :, off + C:off + 2 * C], heads, d, d ** -0.5)
            b.linear(a2.view(tok, C), T[f"{blk}.attn2.out.w"], bias=T[f"{blk}.attn2.out.b"], residual=t, out=t, rows_per_sample=rps)
            n3 = b.layernorm(t, T[f"{blk}.norm3.g"], T[f"{blk}.norm3.b"], 1e-5)
            gg = b.linear(n3, T[f"{blk}.ff1.w"], bias=T[f"{blk}.ff1.b"], geglu=True, block_n=geglu_block_n(8 * C), rows_per_sample=rps)
            b.linear(gg, T[f"{blk}.ff2.w"], bias=T[f"{blk}.ff2.b"], residual=t, out=t, rows_per_sample=rps)
        out = b.linear(t, T[f"{p}.proj_out.w"], bias=T[f"{p}.proj_out.b"], residual=x.view(tok, C), rows_per_sample=rps)
        return out.view(NB, H, W, C)

    def project_context(self, b, ctx):
        """Every cross-attention K/V projection of the network as ONE GEMM over the prompt: [NB*ctx_len, sum 2C]."""
        return b.linear(ctx.view(self.NB * self.ctx_len, ctx.shape[-1]), self.T["ctx_kv.w"], rows_per_sample=self.ctx_len)

    def emit(self, unet_in, ctx, t_explicit=None, step_table=None, step_counter=None, text_embeds=None, time_ids=None,
             eps_out=None, ctx_kv=None):
        """unet_in: bf16 [NB, h*w, cin_pad] (already scaled); ctx: bf16 [NB, ctx_len, D].  Returns eps bf16 [NB, h*w, C_out]
        (written into `eps_out` when given: the pair split lets conv_out store straight into its window slot).
        `ctx_kv`: the cross-attention K/V projections of the prompt when the caller computed them already (they do not
        depend on the timestep: the fused loop projects the prompt once per trajectory, not once per step)."""
        b, T, cfg, s = self.b, self.T, self.cfg, self.s
        NB, h, w = self.NB, self.h, self.w
        eps_r = cfg["norm_eps"]
        c0 = cfg["block_out_channels"][0]
        # --- time embedding ---------------------------------------------------------
        te = b.timestep_embed(NB, c0, t=t_explicit, t_stride=0, step_table=step_table, step_counter=step_counter,
                              flip_sin_to_cos=cfg["flip_sin_to_cos"], freq_shift=cfg["freq_shift"])
        e1 = b.linear(te, T["time_embedding.linear_1.w"], bias=T["time_embedding.linear_1.b"], silu=True, rows_per_sample=1)
        if cfg.get("addition_embed_type") == "text_time":
            # SDXL: emb = time_embedding(t) + add_embedding([text_embeds | embed(time_ids)]); resnets consume silu(emb).
            # emb is kept fp32 and enters the last GEMM as a per-row vector, which the epilogue adds BEFORE the SiLU.
            emb = b.linear(e1, T["time_embedding.linear_2.w"], bias=T["time_embedding.linear_2.b"], out_f32=True, rows_per_sample=1)
            adim = cfg["addition_time_embed_dim"]
            nids = time_ids.shape[1]
            tid = b.timestep_embed(NB * nids, adim, t=time_ids.view(-1), t_stride=1,
                                   flip_sin_to_cos=cfg["flip_sin_to_cos"], freq_shift=cfg["freq_shift"])
            a1 = b.linear(text_embeds, T["add_embedding.linear_1.w"], bias=T["add_embedding.linear_1.b"], silu=True,
                          a2=tid.view(NB, nids * adim), rows_per_sample=1)
            silu_emb = b.linear(a1, T["add_embedding.linear_2.w"], bias=T["add_embedding.linear_2.b"], rowvec=emb,
                                rows_per_group=1, silu=True, rows_per_sample=1)
        else:
            silu_emb = b.linear(e1, T["time_embedding.linear_2.w"], bias=T["time_embedding.linear_2.b"], silu=True, rows_per_sample=1)
        self.temb_all = b.linear(silu_emb, T["temb_all.w"], bias=T["temb_all.b"], out_f32=True, rows_per_sample=1)
        # --- every cross-attention K/V projection of the network in one GEMM --------------------------
        self.ctx_kv = ctx_kv if ctx_kv is not None else self.project_context(b, ctx)
        # --- down path --------------------------------------------------------------------------------
        x = b.conv(unet_in.view(NB, h, w, self.pk.cin_pad), T["conv_in.w"], bias=T["conv_in.b"])
        skips = [x]
        for blk in s["down"]:
            for j, (_, _, cout) in enumerate(blk["resnets"]):
                x = self.resnet(f"{blk['name']}.resnets.{j}", x, None, cout, eps_r)
                if blk["attn"]:
                    x = self.transformer(f"{blk['name']}.attentions.{j}", x, blk["attn"]["heads"], blk["attn"]["layers"])
                skips.append(x)
            if blk["down"]:
                p = f"{blk['name']}.downsamplers.0.conv"
                xs = b.phase_split(x)
                x = b.conv(xs, T[f"{p}.w"], bias=T[f"{p}.b"], taps=L.TAPS_3X3_S2_PAD1, phases=4)
                skips.append(x)
        # --- mid --------------------------------------------------------------------------------------
        m = s["mid"]
        x = self.resnet("mid_block.resnets.0", x, None, m["ch"], eps_r)
        x = self.transformer("mid_block.attentions.0", x, m["attn"]["heads"], m["attn"]["layers"])
        x = self.resnet("mid_block.resnets.1", x, None, m["ch"], eps_r)
        # --- up path ----------------------------------------------------------------------------------
        for blk in s["up"]:
            for j, (_, _, cout) in enumerate(blk["resnets"]):
                x = self.resnet(f"{blk['name']}.resnets.{j}", x, skips.pop(), cout, eps_r)
                if blk["attn"]:
                    x = self.transformer(f"{blk['name']}.attentions.{j}", x, blk["attn"]["heads"], blk["attn"]["layers"])
            if blk["up"]:
                p = f"{blk['name']}.upsamplers.0.conv"
                x = b.upsample2x_conv(x, T[f"{p}.w"], T[f"{p}.wph"], bias=T[f"{p}.b"])
        NBx, H, W, C = x.shape
        hn = b.groupnorm(x.view(NB, H * W, C), T["conv_norm_out.g"], T["conv_norm_out.b"], eps_r, True)
        eps = b.conv(hn.view(NB, H, W, C), T["conv_out.w"], bias=T["conv_out.b"],
                     out=None if eps_out is None else eps_out.view(NB, H, W, eps_out.shape[-1]))
        return eps.view(NB, H * W, eps.shape[-1])


# ------------------------------------------------------------------------------------------------------
# the drop-in object
# ------------------------------------------------------------------------------------------------------
class _Linear1Shim:
    def __init__(self, in_features):
        self.in_features = in_features


class _AddEmbeddingShim:
    def __init__(self, in_features):
        self.linear_1 = _Linear1Shim(in_features)


class B200UNet:
    """Duck-types the members of `model.base` the reference touches."""

    dtype = torch.bfloat16

    def __init__(self, config: Optional[Config] = None, state_dict: Optional[Dict[str, torch.Tensor]] = None,
                 device="cuda", ctx_len: int = 77):
        self.config = config if config is not None else unet_config()
        self.device = torch.device(device)
        self.ctx_len = ctx_len
        self._master = state_dict  # pristine (un-merged) weights, kept on the host for LoRA re-merges
        self._adapters: "OrderedDict[str, dict]" = OrderedDict()
        self._adapter_weights: Dict[str, float] = {}
        self.packed: Optional[PackedUNet] = None
        self.pack_version = 0  # bumped by every re-pack (LoRA load / set / delete): cache key of everything built on `packed`
        self._programs: Dict[tuple, dict] = {}
        if cfg_is_sdxl(self.config):
            self.add_embedding = _AddEmbeddingShim(self.config["projection_class_embeddings_input_dim"])
        if state_dict is not None and self.device.type == "cuda":
            self._repack()

    # ---- weights ------------------------------------------------------------------------------------
    def _repack(self):
        from .lora import merged_state_dict

        sd = merged_state_dict(self._master, self._adapters, self._adapter_weights, prefix="unet")
        self.packed = PackedUNet(self.config, sd, self.device)
        self.pack_version += 1
        self._programs.clear()

    def to(self, device=None, dtype=None):
        if device is not None and torch.device(device) != self.device:
            self.device = torch.device(device)
            if self.device.type != "cuda":
                raise RuntimeError("B200UNet lives on a CUDA device; there is no CPU path")
            self._repack()
        elif self.packed is None and self._master is not None:
            self._repack()
        return self

    # ---- LoRA surface (models/stable_diffusion.py:298-335) -------------------------------------------
    @property
    def peft_config(self):
        return self._adapters

    def add_lora(self, name: str, layers: dict):
        self._adapters[name] = layers
        self._adapter_weights.setdefault(name, 1.0)
        self._repack()

    def set_adapters(self, adapter_names, weights=None):
        names = [adapter_names] if isinstance(adapter_names, str) else list(adapter_names)
        if weights is None:
            weights = [1.0] * len(names)
        if not isinstance(weights, (list, tuple)):
            weights = [weights] * len(names)
        for n in names:
            if n not in self._adapters:
                raise ValueError(f"Adapter name {n} not found")
        self._adapter_weights = {n: 0.0 for n in self._adapters}
        for n, w in zip(names, weights):
            self._adapter_weights[n] = 1.0 if w is None else float(w)
        self._repack()

    def delete_adapters(self, adapter_names):
        names = [adapter_names] if isinstance(adapter_names, str) else list(adapter_names)
        for n in names:
            self._adapters.pop(n, None)
            self._adapter_weights.pop(n, None)
        self._repack()

    # ---- standalone forward (the reference's un-fused call) ------------------------------------------
    def _program(self, NB, h, w, in_dtype, D):
        key = (NB, h, w, in_dtype, D, N.batch_invariant())  # (plans made in the batch-invariant mode differ)
        if key in self._programs:
            return self._programs[key]
        b = L.Builder(self.device)
        pk = self.packed
        st = dict(b=b)
        st["sample"] = b.empty([NB, self.config["in_channels"], h, w], in_dtype)
        st["t"] = b.empty([1], torch.float32)
        st["ctx"] = b.empty([NB, self.ctx_len, D])
        unet_in = b.empty([NB, h * w, pk.cin_pad])
        b.prep_latents(st["sample"], unet_in, rep=1, fixed_scale=1.0)
        low = UNetLowering(b, pk, NB, h, w, self.ctx_len)
        kw = {}
        if cfg_is_sdxl(self.config):
            st["text_embeds"] = b.empty([NB, self.config["projection_class_embeddings_input_dim"] - 6 * self.config["addition_time_embed_dim"]])
            st["time_ids"] = b.empty([NB, 6], torch.float32)
            kw = dict(text_embeds=st["text_embeds"], time_ids=st["time_ids"])
        eps = low.emit(unet_in, st["ctx"], t_explicit=st["t"], **kw)
        st["out"] = b.empty([NB, self.config["out_channels"], h, w], in_dtype)
        b.image_out(eps, self.config["out_channels"], st["out"].view(NB, self.config["out_channels"], h * w))
        st["graph"] = False
        self._programs[key] = st
        return st

    @torch.no_grad()
    def __call__(self, sample, timestep, encoder_hidden_states, cross_attention_kwargs=None, added_cond_kwargs=None,
                 return_dict=False, **unused):
        if self.packed is None:
            raise RuntimeError("B200UNet has no weights on a CUDA device (call .to('cuda') after loading)")
        if cross_attention_kwargs and cross_attention_kwargs.get("scale", None) not in (None, 1.0):
            raise ValueError("cross_attention_kwargs['scale'] != 1 is applied at merge time: use set_adapters(names, weights)")
        NB, _, h, w = sample.shape
        st = self._program(NB, h, w, sample.dtype, encoder_hidden_states.shape[-1])
        stream = torch.cuda.current_stream(self.device)
        st["sample"].copy_(sample)
        st["ctx"].copy_(encoder_hidden_states)
        t = torch.as_tensor(timestep)
        st["t"].copy_(t.reshape(-1)[:1].to(torch.float32))
        if "text_embeds" in st:
            st["text_embeds"].copy_(added_cond_kwargs["text_embeds"])
            st["time_ids"].copy_(added_cond_kwargs["time_ids"])
        st["b"].prog.run(stream.cuda_stream)
        out = st["out"].clone()
        return (out,)

    @property
    def num_launches(self):
        return {k: v["b"].prog.num_launches for k, v in self._programs.items()}


def cfg_is_sdxl(cfg) -> bool:
    return cfg.get("addition_embed_type") == "text_time"
