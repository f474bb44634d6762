"""Generates the golden vectors under tests/golden/ with the oracle (run here, on CPU; committed with its outputs).

    python -m tests.golden.make_golden

The reference itself ships no golden vectors (it has no tests) and its arithmetic lives in diffusers 0.27.2, which is
not installable in this container; these vectors therefore come from the fp32 restatement in oracle/ with fixed seeds.
Small configurations keep the files tiny (< 300 KB) and the CPU test suite fast."""
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SMALL_UNET = dict(block_out_channels=(64, 128, 128), layers_per_block=1, cross_attention_dim=64, attention_head_dim=2,
                  down_block_types=("CrossAttnDownBlock2D", "CrossAttnDownBlock2D", "DownBlock2D"),
                  up_block_types=("UpBlock2D", "CrossAttnUpBlock2D", "CrossAttnUpBlock2D"))
SMALL_VAE = dict(block_out_channels=(32, 64, 64, 64), layers_per_block=1)


def _bf16(sd):
    return {k: v.to(torch.bfloat16).float() for k, v in sd.items()}


def small_unet_state_dict():
    from stablediffusion_b200.config import unet_config
    from stablediffusion_b200.random_init import random_state_dict
    from stablediffusion_b200.unet import unet_param_shapes

    return _bf16(random_state_dict(unet_param_shapes(unet_config(**SMALL_UNET)), seed=0))


def small_vae_state_dict():
    from stablediffusion_b200.config import vae_config
    from stablediffusion_b200.random_init import random_state_dict
    from stablediffusion_b200.vae import vae_param_shapes

    return _bf16(random_state_dict(vae_param_shapes(vae_config(**SMALL_VAE)), seed=1))


def build_small_unet():
    from oracle.unet import UNet2DConditionOracle

    m = UNet2DConditionOracle(**SMALL_UNET).eval()
    m.load_state_dict(small_unet_state_dict(), strict=True)
    return m


def build_small_vae():
    from oracle.vae import AutoencoderKLOracle

    m = AutoencoderKLOracle(**SMALL_VAE).eval()
    m.load_state_dict(small_vae_state_dict(), strict=True)
    return m


def main():
    from oracle import pipeline as OP
    from oracle import schedulers as OS

    torch.manual_seed(0)
    g = torch.Generator().manual_seed(123)
    unet, vae = build_small_unet(), build_small_vae()
    sample = torch.randn(2, 4, 16, 16, generator=g).to(torch.bfloat16).float()
    ctx = torch.randn(2, 77, 64, generator=g).to(torch.bfloat16).float()
    t = torch.tensor(946.4211)
    with torch.no_grad():
        eps = unet(sample, t, ctx)
    torch.save(dict(sample=sample, timestep=t, context=ctx, eps=eps), os.path.join(HERE, "unet_small.pt"))
    z = torch.randn(1, 4, 16, 16, generator=g).to(torch.bfloat16).float()
    with torch.no_grad():
        img = vae.decode(z)
    torch.save(dict(z=z, image=img), os.path.join(HERE, "vae_small.pt"))
    # a short full trajectory per scheduler: latents after 6 CFG steps + the exact timestep tables
    loops = {}
    for name in ("euler", "DDIM", "DPM++ 2M", "DPM++ 2M Karras"):
        s = OS.make_scheduler(name, dict(OS.SD15_SCHEDULER_CONFIG))
        s.set_timesteps(6)
        lat0 = torch.randn(1, 4, 16, 16, generator=torch.Generator().manual_seed(5))
        pe, ne = ctx[:1], ctx[1:]
        out = OP.denoise(unet, s, lat0 * float(s.init_noise_sigma), torch.cat([ne, pe]), s.timesteps, 7.5)
        loops[name] = dict(timesteps=s.timesteps.clone(), sigmas=None if s.sigmas is None else s.sigmas.clone(), lat0=lat0, latents=out)
    torch.save(dict(loops=loops, pe=ctx[:1], ne=ctx[1:]), os.path.join(HERE, "loop_small.pt"))
    tables = {}
    for name, n in (("euler", 20), ("euler", 30), ("euler", 50), ("DDIM", 20), ("DDIM", 50), ("DPM++ 2M", 20), ("DPM++ 2M", 50),
                    ("DPM++ 2M Karras", 25)):
        s = OS.make_scheduler(name, dict(OS.SD15_SCHEDULER_CONFIG))
        s.set_timesteps(n)
        tables[f"{name}/{n}"] = dict(timesteps=s.timesteps.clone(), sigmas=None if s.sigmas is None else s.sigmas.clone(),
                                     init_noise_sigma=float(s.init_noise_sigma))
    sx = OS.EulerDiscreteOracle(**OS.SDXL_SCHEDULER_CONFIG)
    sx.set_timesteps(30)
    tables["sdxl-euler/30"] = dict(timesteps=sx.timesteps.clone(), sigmas=sx.sigmas.clone(), init_noise_sigma=float(sx.init_noise_sigma))
    torch.save(tables, os.path.join(HERE, "scheduler_tables.pt"))
    print("golden vectors written to", HERE)


if __name__ == "__main__":
    main()
