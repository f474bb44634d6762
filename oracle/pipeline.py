"""The reference's denoising loop restated over the oracle modules (test infrastructure).

Follows pipelines/sd_unified_pipeline.py:465-523 line for line: CFG duplicate (:467-469), scale_model_input
(:472), UNet (:475-482), chunk + `guidance_scale * (text - uncond) + uncond` (:484-486), scheduler.step (:489),
latents / scaling_factor (:521), vae.decode (:523)."""
from __future__ import annotations

import torch


def cfg_combine(noise_pred: torch.Tensor, guidance_scale: float) -> torch.Tensor:
    uncond, text = noise_pred.chunk(2)
    return guidance_scale * (text - uncond) + uncond


@torch.no_grad()
def denoise(unet, scheduler, latents, prompt_embeds, timesteps, guidance_scale, do_cfg=True, added_cond_kwargs=None,
            return_eps=False, step_noise=None, inpaint=None):
    """`step_noise[i]`: the standard-normal draw of step i for the ancestral / SDE samplers (the reference draws it
    inside scheduler.step from the global RNG, pipeline :489; tests pass the draws the device used)."""
    eps_trace = []
    for i, t in enumerate(timesteps):
        x_in = torch.cat([latents] * 2) if do_cfg else latents
        x_in = scheduler.scale_model_input(x_in, t)
        noise_pred = unet(x_in, t, prompt_embeds, added_cond_kwargs=added_cond_kwargs)
        if return_eps:
            eps_trace.append(noise_pred.clone())
        if do_cfg:
            noise_pred = cfg_combine(noise_pred, guidance_scale)
        if step_noise is not None:
            latents = scheduler.step(noise_pred, t, latents, step_noise[i])
        else:
            latents = scheduler.step(noise_pred, t, latents)
        if inpaint is not None:  # reference :492-506 (4-channel UNet): keep the unmasked region on the original image
            init_mask, image_latents, noise = inpaint
            proper = image_latents
            if i < len(timesteps) - 1:
                proper = scheduler.add_noise(image_latents, noise, torch.as_tensor([timesteps[i + 1]]))
            latents = (1 - init_mask) * proper + init_mask * latents
    return (latents, eps_trace) if return_eps else latents


@torch.no_grad()
def decode(vae, latents):
    return vae.decode(latents / vae.config["scaling_factor"])
