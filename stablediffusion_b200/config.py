"""Configuration objects with the attribute *and* mapping access the reference relies on
(`unet.config.sample_size`, `vae.config.scaling_factor`, `scheduler.config` round-tripping through
`from_config` — SURVEY §8b).  Field names and defaults are those of the HF `config.json` files the
reference's loader reads (models/stable_diffusion.py:110-138) and that scripts/convert_from_A1111.py:97-203,
490-511 writes."""
from __future__ import annotations

import json
import os


class Config(dict):
    """dict with attribute access (stands in for diffusers' FrozenDict)."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError:
            raise AttributeError(k) from None

    def __setattr__(self, k, v):
        self[k] = v

    def copy(self):
        return Config(self)


UNET_DEFAULTS = dict(
    sample_size=64, in_channels=4, out_channels=4, flip_sin_to_cos=True, freq_shift=0,
    down_block_types=("CrossAttnDownBlock2D", "CrossAttnDownBlock2D", "CrossAttnDownBlock2D", "DownBlock2D"),
    up_block_types=("UpBlock2D", "CrossAttnUpBlock2D", "CrossAttnUpBlock2D", "CrossAttnUpBlock2D"),
    block_out_channels=(320, 640, 1280, 1280), layers_per_block=2, norm_num_groups=32, norm_eps=1e-5,
    cross_attention_dim=768, transformer_layers_per_block=1, attention_head_dim=8, num_attention_heads=None,
    use_linear_projection=False, addition_embed_type=None, addition_time_embed_dim=None,
    projection_class_embeddings_input_dim=None, act_fn="silu", time_cond_proj_dim=None,
)

SDXL_UNET_OVERRIDES = dict(
    sample_size=128, block_out_channels=(320, 640, 1280),
    down_block_types=("DownBlock2D", "CrossAttnDownBlock2D", "CrossAttnDownBlock2D"),
    up_block_types=("CrossAttnUpBlock2D", "CrossAttnUpBlock2D", "UpBlock2D"),
    attention_head_dim=(5, 10, 20), cross_attention_dim=2048, use_linear_projection=True,
    transformer_layers_per_block=(1, 2, 10), addition_embed_type="text_time", addition_time_embed_dim=256,
    projection_class_embeddings_input_dim=2816,
)

VAE_DEFAULTS = dict(
    in_channels=3, out_channels=3, latent_channels=4, block_out_channels=(128, 256, 512, 512), layers_per_block=2,
    norm_num_groups=32, scaling_factor=0.18215, force_upcast=True, sample_size=512, latents_mean=None, latents_std=None,
)

SCHEDULER_DEFAULTS = dict(
    num_train_timesteps=1000, beta_start=0.00085, beta_end=0.012, beta_schedule="scaled_linear", steps_offset=1,
    clip_sample=False, set_alpha_to_one=False, skip_prk_steps=True, prediction_type="epsilon", trained_betas=None,
)
SDXL_SCHEDULER_OVERRIDES = dict(timestep_spacing="leading", interpolation_type="linear", use_karras_sigmas=False,
                                sample_max_value=1.0)


def unet_config(model_type: str = "sd15", **overrides) -> Config:
    c = Config(UNET_DEFAULTS)
    if model_type == "sdxl":
        c.update(SDXL_UNET_OVERRIDES)
    c.update(overrides)
    for k in ("down_block_types", "up_block_types", "block_out_channels"):
        c[k] = tuple(c[k])
    return c


def vae_config(model_type: str = "sd15", **overrides) -> Config:
    c = Config(VAE_DEFAULTS)
    if model_type == "sdxl":
        c["scaling_factor"] = 0.13025
        c["sample_size"] = 1024
    c.update(overrides)
    c["block_out_channels"] = tuple(c["block_out_channels"])
    return c


def scheduler_config(model_type: str = "sd15", **overrides) -> Config:
    c = Config(SCHEDULER_DEFAULTS)
    if model_type == "sdxl":
        c.update(SDXL_SCHEDULER_OVERRIDES)
    c.update(overrides)
    return c


def load_json_config(folder: str, defaults: dict, name: str = "config.json") -> Config:
    with open(os.path.join(folder, name)) as f:
        raw = json.load(f)
    c = Config(defaults)
    c.update({k: v for k, v in raw.items() if not k.startswith("_")})
    return c
