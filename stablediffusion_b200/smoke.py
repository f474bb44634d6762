"""One small invocation of the hot path on cuda:0 (fused loop + VAE decode through the C ABI), checked against the
oracle.  This is one of the few places allowed to import `oracle/` — as the checker, never as the product path."""
from __future__ import annotations

import math

import torch


def run_smoke() -> None:
    if not torch.cuda.is_available():
        raise RuntimeError("smoke() needs a CUDA device: stablediffusion_b200 has no CPU fallback")
    from oracle import pipeline as OP
    from oracle import schedulers as OS
    from oracle.unet import UNet2DConditionOracle
    from oracle.vae import AutoencoderKLOracle

    from .config import unet_config, vae_config
    from .models.stable_diffusion import SDModelWrapper
    from .pipelines.sd_unified_pipeline import StableDiffusionUnifiedPipeline
    from .random_init import random_state_dict
    from .unet import B200UNet, unet_param_shapes
    from .vae import B200VAE, vae_param_shapes

    dev = torch.device("cuda", 0)
    small = dict(block_out_channels=(64, 128, 128), layers_per_block=1, cross_attention_dim=64, attention_head_dim=2,
                 down_block_types=("CrossAttnDownBlock2D", "CrossAttnDownBlock2D", "DownBlock2D"),
                 up_block_types=("UpBlock2D", "CrossAttnUpBlock2D", "CrossAttnUpBlock2D"), sample_size=16)
    vsmall = dict(block_out_channels=(32, 64, 64, 64), layers_per_block=1)
    ucfg, vcfg = unet_config(**small), vae_config(**vsmall)
    rnd = lambda sd: {k: v.to(torch.bfloat16).float() for k, v in sd.items()}
    usd = rnd(random_state_dict(unet_param_shapes(ucfg), seed=0))
    vsd = rnd(random_state_dict(vae_param_shapes(vcfg, True, False), seed=1))
    model = SDModelWrapper(vae=B200VAE(vcfg, vsd, device=dev), base=B200UNet(ucfg, usd, device=dev), device=str(dev))
    g = torch.Generator().manual_seed(7)
    lat = torch.randn(1, 4, 16, 16, generator=g)
    pe = torch.randn(1, 77, 64, generator=g).to(torch.bfloat16)
    ne = torch.randn(1, 77, 64, generator=g).to(torch.bfloat16)
    pipe = StableDiffusionUnifiedPipeline(do_cfg=True, device=str(dev))
    img = pipe(model, prompt_embeds=pe.to(dev), negative_prompt_embeds=ne.to(dev), num_inference_steps=4, guidance_scale=7.5,
               latents=lat.to(dev), height=128, width=128).float().cpu()

    ou = UNet2DConditionOracle(**{k: small[k] for k in small if k != "sample_size"}).eval()
    ou.load_state_dict(usd, strict=True)
    ov = AutoencoderKLOracle(**vsmall).eval()
    ov.load_state_dict(vsd, strict=False)
    sch = OS.make_scheduler("euler", dict(OS.SD15_SCHEDULER_CONFIG))
    sch.set_timesteps(4)
    z = OP.denoise(ou, sch, lat * float(sch.init_noise_sigma), torch.cat([ne, pe]).float(), sch.timesteps, 7.5)
    with torch.no_grad():
        ref = OP.decode(ov, z)
    mse = (img - ref).pow(2).mean().item()
    peak = (ref.max() - ref.min()).item()
    psnr = 10 * math.log10(peak * peak / max(mse, 1e-20))
    print(f"smoke: image {tuple(img.shape)}, PSNR vs oracle {psnr:.1f} dB, launches/step {pipe.last_engine.launches_per_step}")
    if not (torch.isfinite(img).all() and psnr >= 30.0):
        raise AssertionError(f"smoke parity failed: PSNR {psnr:.2f} dB")
