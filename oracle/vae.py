"""fp32 PyTorch restatement of diffusers 0.27.2 `AutoencoderKL` (decoder + encoder), SD1.5 / SDXL topology.

Test infrastructure (see oracle/__init__.py).  Reference call sites: decode at
pipelines/sd_unified_pipeline.py:523, encode at :1017-1041; state-dict key names as produced by
scripts/convert_from_A1111.py:572-677, configuration from :490-511.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from .unet import Downsample2D, ResnetBlock2D, Upsample2D

SD_VAE = dict(in_channels=3, out_channels=3, latent_channels=4, block_out_channels=(128, 256, 512, 512),
              layers_per_block=2, norm_num_groups=32, scaling_factor=0.18215, force_upcast=True, sample_size=512)


class VAEAttention(nn.Module):
    """Single-head attention of the VAE mid block (diffusers `Attention` built from the deprecated AttnBlock:
    GroupNorm(eps 1e-6) -> q,k,v Linear with bias -> softmax(q k^T / sqrt(C)) v -> Linear -> + residual)."""

    def __init__(self, ch, groups=32):
        super().__init__()
        self.group_norm = nn.GroupNorm(groups, ch, eps=1e-6)
        self.to_q = nn.Linear(ch, ch)
        self.to_k = nn.Linear(ch, ch)
        self.to_v = nn.Linear(ch, ch)
        self.to_out = nn.ModuleList([nn.Linear(ch, ch), nn.Dropout(0.0)])

    def forward(self, x):
        B, C, H, W = x.shape
        h = self.group_norm(x).view(B, C, H * W).transpose(1, 2)
        q, k, v = self.to_q(h), self.to_k(h), self.to_v(h)
        o = F.scaled_dot_product_attention(q[:, None], k[:, None], v[:, None])[:, 0]
        o = self.to_out[0](o)
        return o.transpose(1, 2).reshape(B, C, H, W) + x


class VAEMidBlock(nn.Module):
    def __init__(self, ch, groups):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(ch, ch, None, groups, 1e-6) for _ in range(2)])
        self.attentions = nn.ModuleList([VAEAttention(ch, groups)])

    def forward(self, x):
        x = self.resnets[0](x)
        x = self.attentions[0](x)
        return self.resnets[1](x)


class UpDecoderBlock(nn.Module):
    def __init__(self, cin, cout, layers, add_up, groups):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(cin if i == 0 else cout, cout, None, groups, 1e-6) for i in range(layers)])
        self.upsamplers = nn.ModuleList([Upsample2D(cout)]) if add_up else None

    def forward(self, x):
        for r in self.resnets:
            x = r(x)
        if self.upsamplers is not None:
            x = self.upsamplers[0](x)
        return x


class DownEncoderBlock(nn.Module):
    def __init__(self, cin, cout, layers, add_down, groups):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(cin if i == 0 else cout, cout, None, groups, 1e-6) for i in range(layers)])
        self.downsamplers = nn.ModuleList([Downsample2D(cout, padding=0)]) if add_down else None

    def forward(self, x):
        for r in self.resnets:
            x = r(x)
        if self.downsamplers is not None:
            x = self.downsamplers[0](x)
        return x


class Decoder(nn.Module):
    def __init__(self, c):
        super().__init__()
        boc, g = c["block_out_channels"], c["norm_num_groups"]
        self.conv_in = nn.Conv2d(c["latent_channels"], boc[-1], 3, padding=1)
        self.mid_block = VAEMidBlock(boc[-1], g)
        rev = list(reversed(boc))
        self.up_blocks = nn.ModuleList()
        out_ch = rev[0]
        for i in range(len(boc)):
            prev, out_ch = out_ch, rev[i]
            self.up_blocks.append(UpDecoderBlock(prev, out_ch, c["layers_per_block"] + 1, i != len(boc) - 1, g))
        self.conv_norm_out = nn.GroupNorm(g, boc[0], eps=1e-6)
        self.conv_out = nn.Conv2d(boc[0], c["out_channels"], 3, padding=1)

    def forward(self, z):
        x = self.mid_block(self.conv_in(z))
        for blk in self.up_blocks:
            x = blk(x)
        return self.conv_out(F.silu(self.conv_norm_out(x)))


class Encoder(nn.Module):
    def __init__(self, c):
        super().__init__()
        boc, g = c["block_out_channels"], c["norm_num_groups"]
        self.conv_in = nn.Conv2d(c["in_channels"], boc[0], 3, padding=1)
        self.down_blocks = nn.ModuleList()
        out_ch = boc[0]
        for i in range(len(boc)):
            in_ch, out_ch = out_ch, boc[i]
            self.down_blocks.append(DownEncoderBlock(in_ch, out_ch, c["layers_per_block"], i != len(boc) - 1, g))
        self.mid_block = VAEMidBlock(boc[-1], g)
        self.conv_norm_out = nn.GroupNorm(g, boc[-1], eps=1e-6)
        self.conv_out = nn.Conv2d(boc[-1], 2 * c["latent_channels"], 3, padding=1)

    def forward(self, x):
        x = self.conv_in(x)
        for blk in self.down_blocks:
            x = blk(x)
        x = self.mid_block(x)
        return self.conv_out(F.silu(self.conv_norm_out(x)))


class DiagonalGaussian:
    """diffusers `DiagonalGaussianDistribution`: sample = mean + exp(0.5*clamp(logvar,-30,20)) * randn(generator)."""

    def __init__(self, moments):
        self.mean, logvar = moments.chunk(2, dim=1)
        self.logvar = logvar.clamp(-30.0, 20.0)
        self.std = torch.exp(0.5 * self.logvar)

    def sample(self, generator=None):
        noise = torch.randn(self.mean.shape, generator=generator, device=self.mean.device, dtype=self.mean.dtype)
        return self.mean + self.std * noise

    def mode(self):
        return self.mean


class AutoencoderKLOracle(nn.Module):
    def __init__(self, **cfg):
        super().__init__()
        c = dict(SD_VAE)
        c.update(cfg)
        self.config = c
        self.encoder = Encoder(c)
        self.decoder = Decoder(c)
        self.quant_conv = nn.Conv2d(2 * c["latent_channels"], 2 * c["latent_channels"], 1)
        self.post_quant_conv = nn.Conv2d(c["latent_channels"], c["latent_channels"], 1)

    def decode(self, z):
        return self.decoder(self.post_quant_conv(z))

    def encode(self, x) -> DiagonalGaussian:
        return DiagonalGaussian(self.quant_conv(self.encoder(x)))
