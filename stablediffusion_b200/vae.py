"""B200 VAE: the replacement for `model.vae` (diffusers `AutoencoderKL`) on the reference's hot path —
decode at pipelines/sd_unified_pipeline.py:523 (after the latent un-scale of :511-521), encode for img2img at
:1017-1041; loader models/stable_diffusion.py:110-116; key names scripts/convert_from_A1111.py:572-677.

Same op library as the UNet (implicit-GEMM convs, fused GroupNorm+SiLU, residual epilogues).  The mid-block
attention is single-head with d = 512: its score matrix is materialised (q k^T GEMM -> row softmax -> P V GEMM,
with V produced already transposed by swapping the GEMM operands), which costs 3 launches per image and
~1.5 % of the decoder's FLOPs."""
from __future__ import annotations

from collections import OrderedDict
from typing import Dict, Optional, Tuple

import torch

from . import _native as N
from . import lowering as L
from .config import Config, vae_config
from .packing import conv_weight_to_gemm, upsample_conv_phase_weights


def _resnet_shapes(p, cin, cout, sh):
    sh[f"{p}.norm1.weight"] = (cin,); sh[f"{p}.norm1.bias"] = (cin,)
    sh[f"{p}.conv1.weight"] = (cout, cin, 3, 3); sh[f"{p}.conv1.bias"] = (cout,)
    sh[f"{p}.norm2.weight"] = (cout,); sh[f"{p}.norm2.bias"] = (cout,)
    sh[f"{p}.conv2.weight"] = (cout, cout, 3, 3); sh[f"{p}.conv2.bias"] = (cout,)
    if cin != cout:
        sh[f"{p}.conv_shortcut.weight"] = (cout, cin, 1, 1); sh[f"{p}.conv_shortcut.bias"] = (cout,)


def _mid_shapes(p, ch, sh):
    _resnet_shapes(f"{p}.resnets.0", ch, ch, sh)
    a = f"{p}.attentions.0"
    sh[f"{a}.group_norm.weight"] = (ch,); sh[f"{a}.group_norm.bias"] = (ch,)
    for n in ("to_q", "to_k", "to_v", "to_out.0"):
        sh[f"{a}.{n}.weight"] = (ch, ch); sh[f"{a}.{n}.bias"] = (ch,)
    _resnet_shapes(f"{p}.resnets.1", ch, ch, sh)


def vae_structure(cfg: Config) -> dict:
    boc = tuple(cfg["block_out_channels"])
    lpb = cfg["layers_per_block"]
    dec, out_ch = [], boc[-1]
    for i, ch in enumerate(reversed(boc)):
        prev, out_ch = out_ch, ch
        dec.append(dict(name=f"decoder.up_blocks.{i}", resnets=[(prev if j == 0 else out_ch, out_ch) for j in range(lpb + 1)],
                        up=(i != len(boc) - 1), ch=out_ch))
    enc, out_ch = [], boc[0]
    for i, ch in enumerate(boc):
        in_ch, out_ch = out_ch, ch
        enc.append(dict(name=f"encoder.down_blocks.{i}", resnets=[(in_ch if j == 0 else out_ch, out_ch) for j in range(lpb)],
                        down=(i != len(boc) - 1), ch=out_ch))
    return dict(dec=dec, enc=enc, mid=boc[-1])


def vae_param_shapes(cfg: Config, decoder: bool = True, encoder: bool = True) -> "OrderedDict[str, Tuple[int, ...]]":
    s = vae_structure(cfg)
    boc, lc = cfg["block_out_channels"], cfg["latent_channels"]
    sh: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict()
    if encoder:
        sh["encoder.conv_in.weight"] = (boc[0], cfg["in_channels"], 3, 3); sh["encoder.conv_in.bias"] = (boc[0],)
        for blk in s["enc"]:
            for j, (cin, cout) in enumerate(blk["resnets"]):
                _resnet_shapes(f"{blk['name']}.resnets.{j}", cin, cout, sh)
            if blk["down"]:
                sh[f"{blk['name']}.downsamplers.0.conv.weight"] = (blk["ch"], blk["ch"], 3, 3)
                sh[f"{blk['name']}.downsamplers.0.conv.bias"] = (blk["ch"],)
        _mid_shapes("encoder.mid_block", s["mid"], sh)
        sh["encoder.conv_norm_out.weight"] = (boc[-1],); sh["encoder.conv_norm_out.bias"] = (boc[-1],)
        sh["encoder.conv_out.weight"] = (2 * lc, boc[-1], 3, 3); sh["encoder.conv_out.bias"] = (2 * lc,)
        sh["quant_conv.weight"] = (2 * lc, 2 * lc, 1, 1); sh["quant_conv.bias"] = (2 * lc,)
    if decoder:
        sh["post_quant_conv.weight"] = (lc, lc, 1, 1); sh["post_quant_conv.bias"] = (lc,)
        sh["decoder.conv_in.weight"] = (boc[-1], lc, 3, 3); sh["decoder.conv_in.bias"] = (boc[-1],)
        _mid_shapes("decoder.mid_block", s["mid"], sh)
        for blk in s["dec"]:
            for j, (cin, cout) in enumerate(blk["resnets"]):
                _resnet_shapes(f"{blk['name']}.resnets.{j}", cin, cout, sh)
            if blk["up"]:
                sh[f"{blk['name']}.upsamplers.0.conv.weight"] = (blk["ch"], blk["ch"], 3, 3)
                sh[f"{blk['name']}.upsamplers.0.conv.bias"] = (blk["ch"],)
        sh["decoder.conv_norm_out.weight"] = (boc[0],); sh["decoder.conv_norm_out.bias"] = (boc[0],)
        sh["decoder.conv_out.weight"] = (cfg["out_channels"], boc[0], 3, 3); sh["decoder.conv_out.bias"] = (cfg["out_channels"],)
    return sh


def _pad8(n):
    return (n + 7) // 8 * 8


class PackedVAE:
    def __init__(self, cfg: Config, sd: Dict[str, torch.Tensor], device, decoder=True, encoder=True):
        self.cfg, self.device = cfg, device
        self.s = vae_structure(cfg)
        self.has_encoder = encoder and "encoder.conv_in.weight" in sd
        shapes = vae_param_shapes(cfg, decoder, self.has_encoder)
        for k, shp in shapes.items():
            if k not in sd:
                raise ValueError(f"VAE state dict is missing {k}")
            got = tuple(sd[k].shape)
            if got != tuple(shp) and not (len(shp) == 2 and got == tuple(shp) + (1, 1)):  # old checkpoints store attn as 1x1 conv
                raise ValueError(f"VAE tensor {k} has shape {got}, expected {tuple(shp)}")
        self._sd = sd
        self.t: Dict[str, torch.Tensor] = {}
        self._pack(decoder)
        self._sd = None

    def _w(self, n):
        return self._sd[n].to(device=self.device, dtype=torch.float32)

    def _bf(self, t):
        return t.to(torch.bfloat16).contiguous()

    def _f32(self, t):
        return t.to(torch.float32).contiguous()

    def _mat(self, n):
        w = self._w(n)
        return w.reshape(w.shape[0], w.shape[1])

    def _conv_small_cin(self, name):
        """3x3 conv whose C_in (3 or 4) is padded to 8 so a pixel row is 16 bytes."""
        w = self._w(f"{name}.weight")
        co, ci = w.shape[:2]
        wp = torch.zeros(co, 3, 3, _pad8(ci), device=self.device)
        wp[..., :ci] = w.permute(0, 2, 3, 1)
        self.t[f"{name}.w"], self.t[f"{name}.b"] = self._bf(wp.reshape(co, -1)), self._f32(self._w(f"{name}.bias"))

    def _conv(self, name):
        self.t[f"{name}.w"] = self._bf(conv_weight_to_gemm(self._w(f"{name}.weight")))
        self.t[f"{name}.b"] = self._f32(self._w(f"{name}.bias"))

    def _pointwise_padded(self, name):
        """1x1 conv between tiny channel counts (4->4, 8->8): pad both sides to a multiple of 8."""
        w = self._mat(f"{name}.weight")
        co, ci = w.shape
        wp = torch.zeros(_pad8(co), _pad8(ci), device=self.device)
        wp[:co, :ci] = w
        bp = torch.zeros(_pad8(co), device=self.device)
        bp[:co] = self._w(f"{name}.bias")
        self.t[f"{name}.w"], self.t[f"{name}.b"] = self._bf(wp), self._f32(bp)

    def _resnet(self, p, shortcut):
        T = self.t
        for n in ("norm1", "norm2"):
            T[f"{p}.{n}.g"], T[f"{p}.{n}.b"] = self._f32(self._w(f"{p}.{n}.weight")), self._f32(self._w(f"{p}.{n}.bias"))
        self._conv(f"{p}.conv1")
        w2 = conv_weight_to_gemm(self._w(f"{p}.conv2.weight"))
        b2 = self._w(f"{p}.conv2.bias")
        if shortcut:
            w2 = torch.cat([w2, self._mat(f"{p}.conv_shortcut.weight")], dim=1)
            b2 = b2 + self._w(f"{p}.conv_shortcut.bias")
        T[f"{p}.conv2.w"], T[f"{p}.conv2.b"] = self._bf(w2), self._f32(b2)

    def _mid(self, p):
        T = self.t
        self._resnet(f"{p}.resnets.0", False)
        self._resnet(f"{p}.resnets.1", False)
        a = f"{p}.attentions.0"
        T[f"{a}.norm.g"], T[f"{a}.norm.b"] = self._f32(self._w(f"{a}.group_norm.weight")), self._f32(self._w(f"{a}.group_norm.bias"))
        T[f"{a}.qk.w"] = self._bf(torch.cat([self._mat(f"{a}.to_q.weight"), self._mat(f"{a}.to_k.weight")], 0))
        T[f"{a}.qk.b"] = self._f32(torch.cat([self._w(f"{a}.to_q.bias"), self._w(f"{a}.to_k.bias")], 0))
        T[f"{a}.v.w"], T[f"{a}.v.b"] = self._bf(self._mat(f"{a}.to_v.weight")), self._f32(self._w(f"{a}.to_v.bias"))
        T[f"{a}.out.w"], T[f"{a}.out.b"] = self._bf(self._mat(f"{a}.to_out.0.weight")), self._f32(self._w(f"{a}.to_out.0.bias"))

    def _pack(self, decoder):
        s = self.s
        if decoder:
            self._pointwise_padded("post_quant_conv")
            self._conv_small_cin("decoder.conv_in")
            self._mid("decoder.mid_block")
            for blk in s["dec"]:
                for j, (cin, cout) in enumerate(blk["resnets"]):
                    self._resnet(f"{blk['name']}.resnets.{j}", cin != cout)
                if blk["up"]:
                    self._conv(f"{blk['name']}.upsamplers.0.conv")
                    self.t[f"{blk['name']}.upsamplers.0.conv.wph"] = self._bf(
                        upsample_conv_phase_weights(self._w(f"{blk['name']}.upsamplers.0.conv.weight")))
            self.t["decoder.conv_norm_out.g"] = self._f32(self._w("decoder.conv_norm_out.weight"))
            self.t["decoder.conv_norm_out.b"] = self._f32(self._w("decoder.conv_norm_out.bias"))
            self._conv("decoder.conv_out")
        if self.has_encoder:
            self._conv_small_cin("encoder.conv_in")
            for blk in s["enc"]:
                for j, (cin, cout) in enumerate(blk["resnets"]):
                    self._resnet(f"{blk['name']}.resnets.{j}", cin != cout)
                if blk["down"]:
                    self._conv(f"{blk['name']}.downsamplers.0.conv")
            self._mid("encoder.mid_block")
            self.t["encoder.conv_norm_out.g"] = self._f32(self._w("encoder.conv_norm_out.weight"))
            self.t["encoder.conv_norm_out.b"] = self._f32(self._w("encoder.conv_norm_out.bias"))
            self._conv("encoder.conv_out")
            self._pointwise_padded("quant_conv")


class VAELowering:
    def __init__(self, b: L.Builder, pk: PackedVAE):
        self.b, self.pk, self.T = b, pk, pk.t

    def resnet(self, p, x, cout):
        b, T = self.b, self.T
        NB, H, W, C = x.shape
        h1 = b.groupnorm(x.view(NB, H * W, C), T[f"{p}.norm1.g"], T[f"{p}.norm1.b"], 1e-6, True)
        h2 = b.conv(h1.view(NB, H, W, C), T[f"{p}.conv1.w"], bias=T[f"{p}.conv1.b"])
        h3 = b.groupnorm(h2.view(NB, H * W, cout), T[f"{p}.norm2.g"], T[f"{p}.norm2.b"], 1e-6, True).view(NB, H, W, cout)
        if C != cout:
            return b.conv(h3, T[f"{p}.conv2.w"], bias=T[f"{p}.conv2.b"], extra=[x])
        return b.conv(h3, T[f"{p}.conv2.w"], bias=T[f"{p}.conv2.b"], residual=x)

    def attention(self, a, x):
        """Single head, d = C: softmax(q k^T / sqrt(C)) v with the score matrix materialised per image."""
        b, T = self.b, self.T
        NB, H, W, C = x.shape
        HW = H * W
        hn = b.groupnorm(x.view(NB, HW, C), T[f"{a}.norm.g"], T[f"{a}.norm.b"], 1e-6, False)
        qk = b.linear(hn.view(NB * HW, C), T[f"{a}.qk.w"], bias=T[f"{a}.qk.b"], rows_per_sample=HW).view(NB, HW, 2 * C)
        o = b.empty([NB, HW, C])
        S = b.empty([HW, HW])
        vT = b.empty([C, HW])
        for n in range(NB):
            # V^T [C, HW] = Wv [C, C] . hn_n[HW, C]^T (+ bias per row): operands swapped so PV needs no transpose
            b.linear(T[f"{a}.v.w"], hn[n], out=vT, bias=T[f"{a}.v.b"], bias_row=True)
            b.linear(qk[n, :, :C], qk[n, :, C:], out=S)
            b.softmax_rows(S, C ** -0.5, out=S)
            b.linear(S, vT, out=o[n])
        out = b.linear(o.view(NB * HW, C), T[f"{a}.out.w"], bias=T[f"{a}.out.b"], residual=x.view(NB * HW, C), rows_per_sample=HW)
        return out.view(NB, H, W, C)

    def mid(self, p, x):
        ch = x.shape[3]
        x = self.resnet(f"{p}.resnets.0", x, ch)
        x = self.attention(f"{p}.attentions.0", x)
        return self.resnet(f"{p}.resnets.1", x, ch)

    def decode(self, z_in, NB, h, w):
        """z_in: bf16 [NB, h*w, 8] latents (already divided by scaling_factor).  Returns bf16 [NB, 8h*8w, C_out]."""
        b, T, s = self.b, self.T, self.pk.s
        z = b.linear(z_in.view(NB * h * w, z_in.shape[-1]), T["post_quant_conv.w"], bias=T["post_quant_conv.b"], rows_per_sample=h * w)
        x = b.conv(z.view(NB, h, w, z.shape[-1]), T["decoder.conv_in.w"], bias=T["decoder.conv_in.b"])
        x = self.mid("decoder.mid_block", x)
        for blk in s["dec"]:
            for j, (_, cout) in enumerate(blk["resnets"]):
                x = self.resnet(f"{blk['name']}.resnets.{j}", x, cout)
            if blk["up"]:
                p = f"{blk['name']}.upsamplers.0.conv"
                x = b.upsample2x_conv(x, T[f"{p}.w"], T[f"{p}.wph"], bias=T[f"{p}.b"])
        NB, H, W, C = x.shape
        hn = b.groupnorm(x.view(NB, H * W, C), T["decoder.conv_norm_out.g"], T["decoder.conv_norm_out.b"], 1e-6, True)
        img = b.conv(hn.view(NB, H, W, C), T["decoder.conv_out.w"], bias=T["decoder.conv_out.b"])
        return img.view(NB, H * W, img.shape[-1])

    def encode(self, x_in, NB, H, W):
        """x_in: bf16 [NB, H*W, 8] image (3 valid channels).  Returns bf16 [NB, (H/8)*(W/8), 8] moments (mean | logvar)."""
        b, T, s = self.b, self.T, self.pk.s
        x = b.conv(x_in.view(NB, H, W, x_in.shape[-1]), T["encoder.conv_in.w"], bias=T["encoder.conv_in.b"])
        for blk in s["enc"]:
            for j, (_, cout) in enumerate(blk["resnets"]):
                x = self.resnet(f"{blk['name']}.resnets.{j}", x, cout)
            if blk["down"]:
                p = f"{blk['name']}.downsamplers.0.conv"
                x = b.conv(b.phase_split(x), T[f"{p}.w"], bias=T[f"{p}.b"], taps=L.TAPS_3X3_S2_PAD0, phases=4)
        x = self.mid("encoder.mid_block", x)
        NB, h, w, C = x.shape
        hn = b.groupnorm(x.view(NB, h * w, C), T["encoder.conv_norm_out.g"], T["encoder.conv_norm_out.b"], 1e-6, True)
        m = b.conv(hn.view(NB, h, w, C), T["encoder.conv_out.w"], bias=T["encoder.conv_out.b"])
        q = b.linear(m.view(NB * h * w, m.shape[-1]), T["quant_conv.w"], bias=T["quant_conv.b"], rows_per_sample=h * w)
        return q.view(NB, h * w, q.shape[-1])


class _LatentDist:
    """`vae.encode(x).latent_dist` (diffusers DiagonalGaussianDistribution).  `sample(generator)` — the call the pipeline
    makes (:98-106, :1027) — is ONE library kernel (OP_GAUSS_SAMPLE: mean + exp(0.5*clamp(logvar,-30,20)) * noise in fp32
    from the encoder's fp32 moments); torch only draws the standard-normal noise from the caller's generator, in the
    order the reference draws it (App. D7).  `mean` / `logvar` / `std` are inspection accessors, not on the path."""

    def __init__(self, moments_nchw: torch.Tensor, dtype: torch.dtype):
        self._moments = moments_nchw  # fp32 [B, 2C, h, w]
        self._dtype = dtype
        self._prog = None

    @property
    def mean(self):
        return self._moments.chunk(2, dim=1)[0].to(self._dtype)

    @property
    def logvar(self):
        return self._moments.chunk(2, dim=1)[1].clamp(-30.0, 20.0).to(self._dtype)

    @property
    def std(self):
        return torch.exp(0.5 * self._moments.chunk(2, dim=1)[1].clamp(-30.0, 20.0)).to(self._dtype)

    def sample(self, generator=None):
        m = self._moments
        B, C2, h, w = m.shape
        noise = torch.randn((B, C2 // 2, h, w), generator=generator, device=m.device, dtype=self._dtype)
        if self._prog is None:
            b = L.Builder(m.device, workspace_bytes=16)
            self._noise = b.empty(noise.shape, self._dtype)
            self._out = b.empty(noise.shape, self._dtype)
            b.gauss_sample(m, self._noise, self._out)
            self._prog = b
        self._noise.copy_(noise)
        self._prog.prog.run(torch.cuda.current_stream(m.device).cuda_stream)
        return self._out.clone()

    def mode(self):
        return self.mean


class _EncodeOut:
    def __init__(self, dist):
        self.latent_dist = dist


class B200VAE:
    """Duck-types the members of `model.vae` the reference touches."""

    def __init__(self, config: Optional[Config] = None, state_dict=None, device="cuda"):
        self.config = config if config is not None else vae_config()
        self.device = torch.device(device)
        self.dtype = torch.bfloat16
        self._master = state_dict
        self.packed: Optional[PackedVAE] = None
        self._programs: Dict[tuple, dict] = {}
        if state_dict is not None and self.device.type == "cuda":
            self.packed = PackedVAE(self.config, state_dict, self.device)

    def to(self, device=None, dtype=None):
        # `vae.to(dtype=float32)` (force_upcast, pipeline :1020-1036) is accepted and ignored: the kernels compute
        # in bf16 with fp32 accumulation regardless, which cannot overflow the way the fp16 SDXL VAE does.
        if device is not None and not isinstance(device, torch.dtype) and torch.device(device) != self.device:
            self.device = torch.device(device)
            if self.device.type != "cuda":
                raise RuntimeError("B200VAE lives on a CUDA device; there is no CPU path")
            self.packed = None
            self._programs.clear()
        if self.packed is None and self._master is not None and self.device.type == "cuda":
            self.packed = PackedVAE(self.config, self._master, self.device)
        return self

    def _decode_program(self, NB, h, w, in_dtype, scale=1.0):
        key = ("dec", NB, h, w, in_dtype, float(scale), N.batch_invariant())
        if key in self._programs:
            return self._programs[key]
        b = L.Builder(self.device)
        lc = self.config["latent_channels"]
        st = dict(b=b)
        st["z"] = b.empty([NB, lc, h, w], in_dtype)
        z_in = b.empty([NB, h * w, _pad8(lc)])
        b.prep_latents(st["z"], z_in, rep=1, fixed_scale=float(scale))
        img = VAELowering(b, self.packed).decode(z_in, NB, h, w)
        co = self.config["out_channels"]
        st["out"] = b.empty([NB, co, 8 * h, 8 * w], in_dtype)
        b.image_out(img, co, st["out"].view(NB, co, 64 * h * w))
        self._programs[key] = st
        return st

    # Decoder activations grow to 128 channels at full resolution (64 MiB per image at 512^2, 256 MiB at 1024^2, times the
    # ~60 intermediates of the program): a batch is decoded in chunks of at most this many output pixels per program run
    # (16 images at 512^2, 4 at 1024^2: ~60 GB of intermediates), so any batch size fits the 180 GB of a B200 and reuses ONE planned program.
    MAX_PIXELS_PER_RUN = 4 * 1024 * 1024

    @torch.no_grad()
    def decode(self, z, return_dict=False, scale: float = 1.0, **unused):
        """`scale` multiplies z first (the pipeline passes 1/scaling_factor so the un-scale of reference :521 is fused)."""
        if self.packed is None:
            raise RuntimeError("B200VAE has no weights on a CUDA device")
        NB, _, h, w = z.shape
        chunk = max(1, min(NB, self.MAX_PIXELS_PER_RUN // (64 * h * w)))
        stream = torch.cuda.current_stream(self.device).cuda_stream
        if chunk >= NB:
            st = self._decode_program(NB, h, w, z.dtype, scale)
            st["z"].copy_(z)
            st["b"].prog.run(stream)
            return (st["out"].clone(),)
        outs = []
        for lo in range(0, NB, chunk):
            n = min(chunk, NB - lo)
            st = self._decode_program(n, h, w, z.dtype, scale)
            st["z"].copy_(z[lo:lo + n])
            st["b"].prog.run(stream)
            outs.append(st["out"].clone())
        return (torch.cat(outs, dim=0),)

    def _encode_program(self, NB, H, W, in_dtype):
        key = ("enc", NB, H, W, in_dtype, N.batch_invariant())
        if key in self._programs:
            return self._programs[key]
        if not self.packed.has_encoder:
            raise RuntimeError("this VAE was loaded without encoder weights")
        b = L.Builder(self.device)
        ci, lc = self.config["in_channels"], self.config["latent_channels"]
        st = dict(b=b)
        st["x"] = b.empty([NB, ci, H, W], in_dtype)
        x_in = b.empty([NB, H * W, _pad8(ci)])
        b.prep_latents(st["x"], x_in, rep=1, fixed_scale=1.0)
        mom = VAELowering(b, self.packed).encode(x_in, NB, H, W)
        st["out"] = b.empty([NB, 2 * lc, H // 8, W // 8], torch.float32)
        b.image_out(mom, 2 * lc, st["out"].view(NB, 2 * lc, (H // 8) * (W // 8)))
        self._programs[key] = st
        return st

    @torch.no_grad()
    def encode(self, x, return_dict=True, **unused):
        NB, _, H, W = x.shape
        st = self._encode_program(NB, H, W, x.dtype)
        st["x"].copy_(x)
        st["b"].prog.run(torch.cuda.current_stream(self.device).cuda_stream)
        return _EncodeOut(_LatentDist(st["out"].clone(), x.dtype))
