"""CPU oracle for the denoising hot path — TEST INFRASTRUCTURE, not product code.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` legs may import
this package; nothing under `stablediffusion_b200/` does (tests/test_abi.py::test_product_never_imports_the_oracle enforces it).

PARITY UNPINNED.  The reference (GrafikXxxxxxxYyyyyyyyyyy/StableDiffusion) owns no arithmetic on this path: every tensor op is
executed by the un-vendored third-party package `diffusers==0.27.2` (requirements.txt:39) on
`torch==2.2.2` (requirements.txt:175), neither of which is present in /root/reference or installable here
(no network, no wheel).  The reference also has no tests, golden vectors or fixtures.
This package therefore restates the *published* algorithm of those diffusers 0.27.2 modules in plain fp32
PyTorch, anchored on the reference's own call sites:

  UNet forward        pipelines/sd_unified_pipeline.py:475-482   -> oracle.unet.UNet2DConditionOracle
  scale_model_input   pipelines/sd_unified_pipeline.py:472       -> oracle.schedulers.*.scale_model_input
  CFG combine         pipelines/sd_unified_pipeline.py:484-486   -> oracle.pipeline.cfg_combine
  scheduler.step      pipelines/sd_unified_pipeline.py:489       -> oracle.schedulers.*.step
  latent un-scale     pipelines/sd_unified_pipeline.py:511-521   -> oracle.pipeline.denoise_and_decode
  VAE decode          pipelines/sd_unified_pipeline.py:523       -> oracle.vae.AutoencoderKLOracle.decode
  VAE encode (img2img) pipelines/sd_unified_pipeline.py:1017-1041 -> oracle.vae.AutoencoderKLOracle.encode
  network topology    scripts/convert_from_A1111.py:97-203,240-485,490-511,572-677 (state-dict key names)
  scheduler configs   scripts/convert_from_A1111.py:947-982; models/stable_diffusion.py:135-138,199-227
  LoRA semantics      models/stable_diffusion.py:230-309; pipelines/train_lora_pipeline.py:247-266,497-528

What pins it instead: (1) analytic parameter counts that reproduce the published totals exactly
(859,520,964 / 2,567,463,684 / 49,490,179 / 34,163,592 — tests/test_oracle.py); (2) state-dict key names
identical to the converter's output, so real checkpoints load strictly; (3) well-known scheduler constants
(sigma_max 14.6146, DDIM-50 table 981..1, alphas_cumprod end points); (4) module-level identities
(attention == softmax form, GEGLU == chunked gelu, LoRA merged == unmerged).  If a GPU box ever carries
`diffusers`, tests/test_oracle.py::test_against_installed_diffusers cross-checks the restatement directly.
"""
