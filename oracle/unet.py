"""fp32 PyTorch restatement of diffusers 0.27.2 `UNet2DConditionModel` (SD1.5 and SDXL-base configurations).

Test infrastructure (see oracle/__init__.py).  Follows the reference call site
pipelines/sd_unified_pipeline.py:475-482 and the topology / state-dict key names produced by
scripts/convert_from_A1111.py:97-203 (config derivation) and :240-485 (key names).  Module and parameter
names are identical to diffusers' so a real checkpoint's state dict loads with strict=True.
"""
from __future__ import annotations

import math
from typing import Optional, Sequence

import torch
import torch.nn as nn
import torch.nn.functional as F


def _tuple(v, n):
    return tuple(v) if isinstance(v, (tuple, list)) else (v,) * n


SD15_UNET = dict(
    in_channels=4, out_channels=4, sample_size=64, block_out_channels=(320, 640, 1280, 1280), layers_per_block=2,
    down_block_types=("CrossAttnDownBlock2D", "CrossAttnDownBlock2D", "CrossAttnDownBlock2D", "DownBlock2D"),
    up_block_types=("UpBlock2D", "CrossAttnUpBlock2D", "CrossAttnUpBlock2D", "CrossAttnUpBlock2D"),
    attention_head_dim=8, cross_attention_dim=768, norm_num_groups=32, norm_eps=1e-5,
    use_linear_projection=False, transformer_layers_per_block=1, flip_sin_to_cos=True, freq_shift=0,
    addition_embed_type=None, addition_time_embed_dim=None, projection_class_embeddings_input_dim=None,
)
SDXL_UNET = dict(
    in_channels=4, out_channels=4, sample_size=128, block_out_channels=(320, 640, 1280), layers_per_block=2,
    down_block_types=("DownBlock2D", "CrossAttnDownBlock2D", "CrossAttnDownBlock2D"),
    up_block_types=("CrossAttnUpBlock2D", "CrossAttnUpBlock2D", "UpBlock2D"),
    attention_head_dim=(5, 10, 20), cross_attention_dim=2048, norm_num_groups=32, norm_eps=1e-5,
    use_linear_projection=True, transformer_layers_per_block=(1, 2, 10), flip_sin_to_cos=True, freq_shift=0,
    addition_embed_type="text_time", addition_time_embed_dim=256, projection_class_embeddings_input_dim=2816,
)


def timestep_embedding(timesteps: torch.Tensor, dim: int, flip_sin_to_cos: bool, freq_shift: float) -> torch.Tensor:
    """diffusers `get_timestep_embedding`: fp32, [sin|cos] flipped to [cos|sin]."""
    half = dim // 2
    exponent = -math.log(10000.0) * torch.arange(0, half, dtype=torch.float32, device=timesteps.device)
    exponent = exponent / (half - freq_shift)
    emb = torch.exp(exponent)
    emb = timesteps[:, None].float() * emb[None, :]
    emb = torch.cat([torch.sin(emb), torch.cos(emb)], dim=-1)
    if flip_sin_to_cos:
        emb = torch.cat([emb[:, half:], emb[:, :half]], dim=-1)
    return emb


class TimestepEmbedding(nn.Module):
    def __init__(self, in_dim, dim):
        super().__init__()
        self.linear_1 = nn.Linear(in_dim, dim)
        self.linear_2 = nn.Linear(dim, dim)

    def forward(self, x):
        return self.linear_2(F.silu(self.linear_1(x)))


class ResnetBlock2D(nn.Module):
    def __init__(self, cin, cout, temb_channels: Optional[int], groups=32, eps=1e-5):
        super().__init__()
        self.norm1 = nn.GroupNorm(groups, cin, eps=eps)
        self.conv1 = nn.Conv2d(cin, cout, 3, padding=1)
        self.time_emb_proj = nn.Linear(temb_channels, cout) if temb_channels is not None else None
        self.norm2 = nn.GroupNorm(groups, cout, eps=eps)
        self.conv2 = nn.Conv2d(cout, cout, 3, padding=1)
        self.conv_shortcut = nn.Conv2d(cin, cout, 1) if cin != cout else None

    def forward(self, x, temb=None):
        h = self.conv1(F.silu(self.norm1(x)))
        if self.time_emb_proj is not None:
            h = h + self.time_emb_proj(F.silu(temb))[:, :, None, None]
        h = self.conv2(F.silu(self.norm2(h)))
        if self.conv_shortcut is not None:
            x = self.conv_shortcut(x)
        return x + h  # output_scale_factor = 1


class Attention(nn.Module):
    """diffusers `Attention` with AttnProcessor2_0: softmax(q k^T / sqrt(d)) v, no mask."""

    def __init__(self, query_dim, heads, dim_head, cross_attention_dim=None, bias=False):
        super().__init__()
        inner = heads * dim_head
        self.heads = heads
        self.to_q = nn.Linear(query_dim, inner, bias=bias)
        self.to_k = nn.Linear(cross_attention_dim or query_dim, inner, bias=bias)
        self.to_v = nn.Linear(cross_attention_dim or query_dim, inner, bias=bias)
        self.to_out = nn.ModuleList([nn.Linear(inner, query_dim), nn.Dropout(0.0)])

    def forward(self, x, context=None):
        context = x if context is None else context
        B, Nq, _ = x.shape
        q, k, v = self.to_q(x), self.to_k(context), self.to_v(context)
        d = q.shape[-1] // self.heads

        def split(t):
            return t.view(B, -1, self.heads, d).transpose(1, 2)

        o = F.scaled_dot_product_attention(split(q), split(k), split(v))
        o = o.transpose(1, 2).reshape(B, Nq, self.heads * d)
        return self.to_out[0](o)


class GEGLU(nn.Module):
    def __init__(self, dim_in, dim_out):
        super().__init__()
        self.proj = nn.Linear(dim_in, dim_out * 2)

    def forward(self, x):
        h, g = self.proj(x).chunk(2, dim=-1)
        return h * F.gelu(g)


class FeedForward(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.net = nn.ModuleList([GEGLU(dim, dim * 4), nn.Dropout(0.0), nn.Linear(dim * 4, dim)])

    def forward(self, x):
        return self.net[2](self.net[0](x))


class BasicTransformerBlock(nn.Module):
    def __init__(self, dim, heads, dim_head, cross_attention_dim):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim, eps=1e-5)
        self.attn1 = Attention(dim, heads, dim_head)
        self.norm2 = nn.LayerNorm(dim, eps=1e-5)
        self.attn2 = Attention(dim, heads, dim_head, cross_attention_dim)
        self.norm3 = nn.LayerNorm(dim, eps=1e-5)
        self.ff = FeedForward(dim)

    def forward(self, x, context):
        x = x + self.attn1(self.norm1(x))
        x = x + self.attn2(self.norm2(x), context)
        x = x + self.ff(self.norm3(x))
        return x


class Transformer2DModel(nn.Module):
    def __init__(self, channels, heads, layers, cross_attention_dim, use_linear_projection, groups=32):
        super().__init__()
        self.use_linear_projection = use_linear_projection
        self.norm = nn.GroupNorm(groups, channels, eps=1e-6)
        inner = channels
        if use_linear_projection:
            self.proj_in = nn.Linear(channels, inner)
            self.proj_out = nn.Linear(inner, channels)
        else:
            self.proj_in = nn.Conv2d(channels, inner, 1)
            self.proj_out = nn.Conv2d(inner, channels, 1)
        self.transformer_blocks = nn.ModuleList(
            [BasicTransformerBlock(inner, heads, inner // heads, cross_attention_dim) for _ in range(layers)])

    def forward(self, x, context):
        B, C, H, W = x.shape
        res = x
        h = self.norm(x)
        if not self.use_linear_projection:
            h = self.proj_in(h)
            h = h.permute(0, 2, 3, 1).reshape(B, H * W, C)
        else:
            h = h.permute(0, 2, 3, 1).reshape(B, H * W, C)
            h = self.proj_in(h)
        for blk in self.transformer_blocks:
            h = blk(h, context)
        if not self.use_linear_projection:
            h = h.reshape(B, H, W, C).permute(0, 3, 1, 2)
            h = self.proj_out(h)
        else:
            h = self.proj_out(h)
            h = h.reshape(B, H, W, C).permute(0, 3, 1, 2)
        return h + res


class Downsample2D(nn.Module):
    def __init__(self, channels, padding=1):
        super().__init__()
        self.padding = padding
        self.conv = nn.Conv2d(channels, channels, 3, stride=2, padding=padding)

    def forward(self, x):
        if self.padding == 0:
            x = F.pad(x, (0, 1, 0, 1))
        return self.conv(x)


class Upsample2D(nn.Module):
    def __init__(self, channels):
        super().__init__()
        self.conv = nn.Conv2d(channels, channels, 3, padding=1)

    def forward(self, x):
        return self.conv(F.interpolate(x, scale_factor=2.0, mode="nearest"))


class DownBlock(nn.Module):
    def __init__(self, cin, cout, temb, layers, add_down, attn: Optional[dict], eps, groups):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(cin if i == 0 else cout, cout, temb, groups, eps) for i in range(layers)])
        if attn is not None:
            self.attentions = nn.ModuleList([Transformer2DModel(cout, **attn) for _ in range(layers)])
        else:
            self.attentions = None
        self.downsamplers = nn.ModuleList([Downsample2D(cout)]) if add_down else None

    def forward(self, x, temb, context):
        outs = []
        for i, r in enumerate(self.resnets):
            x = r(x, temb)
            if self.attentions is not None:
                x = self.attentions[i](x, context)
            outs.append(x)
        if self.downsamplers is not None:
            x = self.downsamplers[0](x)
            outs.append(x)
        return x, outs


class MidBlock(nn.Module):
    def __init__(self, ch, temb, attn: dict, eps, groups):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(ch, ch, temb, groups, eps) for _ in range(2)])
        self.attentions = nn.ModuleList([Transformer2DModel(ch, **attn)])

    def forward(self, x, temb, context):
        x = self.resnets[0](x, temb)
        x = self.attentions[0](x, context)
        return self.resnets[1](x, temb)


class UpBlock(nn.Module):
    def __init__(self, cin, cout, prev, temb, layers, add_up, attn: Optional[dict], eps, groups):
        super().__init__()
        rs = []
        for i in range(layers):
            skip = cin if i == layers - 1 else cout
            rin = prev if i == 0 else cout
            rs.append(ResnetBlock2D(rin + skip, cout, temb, groups, eps))
        self.resnets = nn.ModuleList(rs)
        self.attentions = nn.ModuleList([Transformer2DModel(cout, **attn) for _ in range(layers)]) if attn is not None else None
        self.upsamplers = nn.ModuleList([Upsample2D(cout)]) if add_up else None

    def forward(self, x, skips, temb, context):
        for i, r in enumerate(self.resnets):
            x = torch.cat([x, skips.pop()], dim=1)
            x = r(x, temb)
            if self.attentions is not None:
                x = self.attentions[i](x, context)
        if self.upsamplers is not None:
            x = self.upsamplers[0](x)
        return x


class UNet2DConditionOracle(nn.Module):
    def __init__(self, **cfg):
        super().__init__()
        c = dict(SD15_UNET)
        c.update(cfg)
        self.config = c
        boc: Sequence[int] = c["block_out_channels"]
        nb = len(boc)
        heads = _tuple(c["attention_head_dim"], nb)  # historical misnomer: this is the number of heads
        tlayers = _tuple(c["transformer_layers_per_block"], nb)
        groups, eps = c["norm_num_groups"], c["norm_eps"]
        temb = boc[0] * 4
        self.conv_in = nn.Conv2d(c["in_channels"], boc[0], 3, padding=1)
        self.time_embedding = TimestepEmbedding(boc[0], temb)
        if c["addition_embed_type"] == "text_time":
            self.add_embedding = TimestepEmbedding(c["projection_class_embeddings_input_dim"], temb)
        else:
            self.add_embedding = None

        def attn_cfg(i):
            return dict(heads=heads[i], layers=tlayers[i], cross_attention_dim=c["cross_attention_dim"],
                        use_linear_projection=c["use_linear_projection"], groups=groups)

        self.down_blocks = nn.ModuleList()
        out_ch = boc[0]
        for i, t in enumerate(c["down_block_types"]):
            in_ch, out_ch = out_ch, boc[i]
            self.down_blocks.append(DownBlock(in_ch, out_ch, temb, c["layers_per_block"], i != nb - 1,
                                              attn_cfg(i) if t.startswith("CrossAttn") else None, eps, groups))
        self.mid_block = MidBlock(boc[-1], temb, attn_cfg(nb - 1), eps, groups)
        self.up_blocks = nn.ModuleList()
        rev = list(reversed(boc))
        out_ch = rev[0]
        for i, t in enumerate(c["up_block_types"]):
            prev, out_ch = out_ch, rev[i]
            in_ch = rev[min(i + 1, nb - 1)]
            self.up_blocks.append(UpBlock(in_ch, out_ch, prev, temb, c["layers_per_block"] + 1, i != nb - 1,
                                          attn_cfg(nb - 1 - i) if t.startswith("CrossAttn") else None, eps, groups))
        self.conv_norm_out = nn.GroupNorm(groups, boc[0], eps=eps)
        self.conv_out = nn.Conv2d(boc[0], c["out_channels"], 3, padding=1)

    def forward(self, sample, timestep, encoder_hidden_states, added_cond_kwargs=None):
        c = self.config
        B = sample.shape[0]
        t = torch.as_tensor(timestep, device=sample.device)
        if t.dim() == 0:
            t = t[None].expand(B)
        t_emb = timestep_embedding(t, c["block_out_channels"][0], c["flip_sin_to_cos"], c["freq_shift"]).to(sample.dtype)
        emb = self.time_embedding(t_emb)
        if self.add_embedding is not None:
            text_embeds, time_ids = added_cond_kwargs["text_embeds"], added_cond_kwargs["time_ids"]
            tid = timestep_embedding(time_ids.flatten(), c["addition_time_embed_dim"], c["flip_sin_to_cos"], c["freq_shift"])
            tid = tid.reshape(B, -1).to(sample.dtype)
            emb = emb + self.add_embedding(torch.cat([text_embeds, tid], dim=-1))
        x = self.conv_in(sample)
        skips = [x]
        for blk in self.down_blocks:
            x, outs = blk(x, emb, encoder_hidden_states)
            skips.extend(outs)
        x = self.mid_block(x, emb, encoder_hidden_states)
        for blk in self.up_blocks:
            x = blk(x, skips, emb, encoder_hidden_states)
        return self.conv_out(F.silu(self.conv_norm_out(x)))
