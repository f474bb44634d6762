"""Model-level parity on the B200: B200UNet / B200VAE / fused loop (all through the C ABI) vs the fp32 oracle
on identical random-init weights (bf16-rounded), identical latents and context.

Tolerances: per-step UNet eps rel-L2 <= 2e-2 and max-abs <= 0.1*max|eps|; the same quantities of
torch-eager bf16 running the oracle modules on the same GPU are printed beside ours as the yardstick.
"""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

SMALL_UNET = dict(block_out_channels=(64, 128, 128), layers_per_block=1, cross_attention_dim=64, attention_head_dim=2,
                  down_block_types=("CrossAttnDownBlock2D", "CrossAttnDownBlock2D", "DownBlock2D"),
                  up_block_types=("UpBlock2D", "CrossAttnUpBlock2D", "CrossAttnUpBlock2D"), sample_size=16)


def _rel_l2(a, b):
    a, b = a.float(), b.float()
    return ((a - b).pow(2).sum().sqrt() / (b.pow(2).sum().sqrt() + 1e-12)).item()


def _bf16_round_sd(sd):
    return {k: v.to(torch.bfloat16).to(torch.float32) for k, v in sd.items()}


def _make_unet(dev, model_type="sd15", seed=0, **over):
    from oracle.unet import UNet2DConditionOracle
    from stablediffusion_b200.config import unet_config
    from stablediffusion_b200.random_init import random_state_dict
    from stablediffusion_b200.unet import B200UNet, unet_param_shapes

    cfg = unet_config(model_type, **over)
    sd = _bf16_round_sd(random_state_dict(unet_param_shapes(cfg), seed=seed))
    ours = B200UNet(cfg, sd, device=dev)
    okeys = {k: cfg[k] for k in ("in_channels", "out_channels", "block_out_channels", "layers_per_block", "down_block_types",
                                 "up_block_types", "attention_head_dim", "cross_attention_dim", "norm_num_groups", "norm_eps",
                                 "use_linear_projection", "transformer_layers_per_block", "flip_sin_to_cos", "freq_shift",
                                 "addition_embed_type", "addition_time_embed_dim", "projection_class_embeddings_input_dim")}
    oracle = UNet2DConditionOracle(**okeys)
    oracle.load_state_dict(sd, strict=True)
    return cfg, ours, oracle.eval()


def _unet_case(dev, cfg, ours, oracle, B, h, w, t, oracle_on_gpu, max_ratio_vs_eager=None):
    g = torch.Generator().manual_seed(11)
    x = torch.randn(B, 4, h, w, generator=g)
    ctx = torch.randn(B, 77, cfg["cross_attention_dim"], generator=g).to(torch.bfloat16).float()
    added = None
    if cfg.get("addition_embed_type") == "text_time":
        added = dict(text_embeds=torch.randn(B, 1280, generator=g).to(torch.bfloat16).float(),
                     time_ids=torch.tensor([[8.0 * h, 8.0 * w, 0, 0, 8.0 * h, 8.0 * w]]).repeat(B, 1))
    xb = x.to(torch.bfloat16).float()  # the UNet input is bf16 on our path; feed the oracle the same values
    odev = dev if oracle_on_gpu else torch.device("cpu")
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    oracle = oracle.to(odev)
    with torch.no_grad():
        oadd = None if added is None else {k: v.to(odev) for k, v in added.items()}
        ref = oracle(xb.to(odev), torch.tensor(t), ctx.to(odev), added_cond_kwargs=oadd).float().cpu()
        eager = None
        if oracle_on_gpu:
            ob = oracle.to(torch.bfloat16)
            badd = None if added is None else {k: v.to(odev, torch.bfloat16) for k, v in added.items()}
            eager = ob(xb.to(odev, torch.bfloat16), torch.tensor(t), ctx.to(odev, torch.bfloat16), added_cond_kwargs=badd).float().cpu()
            oracle.to(torch.float32)
    dadd = None if added is None else {k: v.to(dev) for k, v in added.items()}
    out = ours(xb.to(dev), torch.tensor(t), ctx.to(dev, torch.bfloat16), added_cond_kwargs=dadd, return_dict=False)[0]
    torch.cuda.synchronize()
    out = out.float().cpu()
    assert torch.isfinite(out).all()
    rl2 = _rel_l2(out, ref)
    mabs = (out - ref).abs().max().item() / ref.abs().max().item()
    msg = f"rel-L2 {rl2:.4g}, max-abs/max {mabs:.4g}"
    if eager is not None:
        msg += f" | torch-eager-bf16 rel-L2 {_rel_l2(eager, ref):.4g}"
    print("UNET PARITY", tuple(x.shape), "t=", t, msg)
    assert rl2 <= 2e-2, msg
    assert mabs <= 0.1, msg
    if max_ratio_vs_eager is not None:  # no worse than 1.5x the error of torch-eager bf16 on the same GPU
        assert eager is not None and rl2 <= max_ratio_vs_eager * _rel_l2(eager, ref), msg
    return rl2


def test_unet_small_vs_oracle_cpu(cuda_device):
    cfg, ours, oracle = _make_unet(cuda_device, **SMALL_UNET)
    _unet_case(cuda_device, cfg, ours, oracle, 2, 16, 16, 946.4211, oracle_on_gpu=False)
    _unet_case(cuda_device, cfg, ours, oracle, 3, 24, 16, 1.0, oracle_on_gpu=False)


def test_unet_sd15_512_vs_oracle(cuda_device):
    cfg, ours, oracle = _make_unet(cuda_device)
    _unet_case(cuda_device, cfg, ours, oracle, 2, 64, 64, 946.4211, oracle_on_gpu=True)
    print("launches per forward:", ours.num_launches)


def test_unet_sd15_768_vs_oracle(cuda_device):
    cfg, ours, oracle = _make_unet(cuda_device)
    _unet_case(cuda_device, cfg, ours, oracle, 2, 96, 96, 501.0, oracle_on_gpu=True)


def test_unet_sdxl_vs_oracle(cuda_device):
    cfg, ours, oracle = _make_unet(cuda_device, "sdxl")
    _unet_case(cuda_device, cfg, ours, oracle, 2, 64, 64, 958.0, oracle_on_gpu=True)


# ---------------------------------------------------------------------------------------------------
# VAE, fused loop, full pipeline
# ---------------------------------------------------------------------------------------------------
def _psnr(a, b, peak=2.0):
    mse = (a.float() - b.float()).pow(2).mean().item()
    return 10.0 * math.log10(peak * peak / max(mse, 1e-20))


def _make_vae(dev, seed=5, encoder=False):
    from oracle.vae import AutoencoderKLOracle
    from stablediffusion_b200.config import vae_config
    from stablediffusion_b200.random_init import random_state_dict
    from stablediffusion_b200.vae import B200VAE, vae_param_shapes

    cfg = vae_config()
    sd = _bf16_round_sd(random_state_dict(vae_param_shapes(cfg, True, True), seed=seed))
    oracle = AutoencoderKLOracle()
    oracle.load_state_dict(sd, strict=True)
    if not encoder:
        sd = {k: v for k, v in sd.items() if not (k.startswith("encoder") or k.startswith("quant_conv"))}
    return cfg, B200VAE(cfg, sd, device=dev), oracle.eval()


@pytest.mark.parametrize("B,h,w", [(1, 32, 32), (2, 64, 64)])
def test_vae_decode_vs_oracle(cuda_device, B, h, w):
    dev = cuda_device
    cfg, ours, oracle = _make_vae(dev)
    z = torch.randn(B, 4, h, w, generator=torch.Generator().manual_seed(3))
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    with torch.no_grad():
        ref = oracle.to(dev).decode(z.to(torch.bfloat16).float().to(dev) ).float().cpu()
    out = ours.decode(z.to(dev), return_dict=False)[0].float().cpu()
    assert out.shape == (B, 3, 8 * h, 8 * w)
    rl2 = _rel_l2(out, ref)
    psnr = _psnr(out, ref, peak=(ref.max() - ref.min()).item())
    print(f"VAE PARITY {B}x{h}x{w}: rel-L2 {rl2:.4g} PSNR {psnr:.1f} dB")
    assert rl2 <= 2e-2 and psnr >= 30.0


def test_vae_encode_vs_oracle(cuda_device):
    dev = cuda_device
    cfg, ours, oracle = _make_vae(dev, encoder=True)
    x = (torch.rand(1, 3, 256, 256, generator=torch.Generator().manual_seed(2)) * 2 - 1).to(torch.bfloat16).float()
    with torch.no_grad():
        ref = oracle.to(dev).encode(x.to(dev))
    d = ours.encode(x.to(dev)).latent_dist
    rl2 = _rel_l2(d.mean.cpu(), ref.mean.cpu())
    print(f"VAE ENCODE PARITY: mean rel-L2 {rl2:.4g}, std rel-L2 {_rel_l2(d.std.cpu(), ref.std.cpu()):.4g}")
    assert rl2 <= 2e-2


@pytest.mark.parametrize("sched_name,steps", [("euler", 12), ("DDIM", 12), ("DPM++ 2M", 12), ("DPM++ 2M Karras", 8),
                                              ("euler_a", 12), ("DPM++ 2M SDE Karras", 12), ("PNDM", 12), ("uni_pc", 12)])
def test_fused_loop_vs_oracle(cuda_device, sched_name, steps):
    """Whole trajectory: B200 fused loop (graph replay) vs the oracle's restatement of the reference loop."""
    from oracle import pipeline as OP
    from oracle import schedulers as OS
    from stablediffusion_b200.models.stable_diffusion import SDModelWrapper
    from stablediffusion_b200.pipelines.sd_unified_pipeline import StableDiffusionUnifiedPipeline
    from stablediffusion_b200.vae import B200VAE

    dev = cuda_device
    cfg, ours, oracle = _make_unet(dev, **SMALL_UNET)
    vcfg, vae, _ = _make_vae(dev)
    model = SDModelWrapper(vae=vae, base=ours, device=str(dev))
    model.set_scheduler(sched_name)
    g = torch.Generator().manual_seed(21)
    B = 2
    lat = torch.randn(B, 4, 16, 16, generator=g)
    pe = torch.randn(B, 77, 64, generator=g).to(torch.bfloat16)
    ne = torch.randn(B, 77, 64, generator=g).to(torch.bfloat16)
    pipe = StableDiffusionUnifiedPipeline(do_cfg=True, device=str(dev), output_type="latents")
    step_noise = None
    if sched_name in OS.STOCHASTIC:  # same device draws for both sides (the engine draws all steps in one randn call)
        pipe._step_generator = torch.Generator(device=dev).manual_seed(77)
        step_noise = torch.randn((steps, B, 4, 16 * 16), generator=torch.Generator(device=dev).manual_seed(77), device=dev,
                                 dtype=torch.float32).view(steps, B, 4, 16, 16).cpu()
    out = pipe(model, prompt_embeds=pe.to(dev), negative_prompt_embeds=ne.to(dev), num_inference_steps=steps,
               guidance_scale=7.5, latents=lat.to(dev), height=128, width=128).float().cpu()
    osched = OS.make_scheduler(sched_name, dict(OS.SD15_SCHEDULER_CONFIG))
    osched.set_timesteps(steps)
    assert torch.equal(osched.timesteps.cpu(), model.scheduler.timesteps.cpu()), "timestep tables differ"
    x0 = lat * float(osched.init_noise_sigma)
    ref = OP.denoise(oracle.cpu(), osched, x0, torch.cat([ne, pe]).float(), osched.timesteps, 7.5, step_noise=step_noise)
    rl2 = _rel_l2(out, ref)
    print(f"LOOP PARITY {sched_name} x{steps}: rel-L2 {rl2:.4g}")
    assert torch.isfinite(out).all() and rl2 <= 3.3e-2   # measured 1.8-2.2e-2 (round 1): 1.5x head-room


@pytest.mark.parametrize("sched_name", ["euler", "PNDM", "euler_a", "DPM++ 2M SDE Karras", "uni_pc"])
def test_unfused_reference_call_pattern_matches_fused(cuda_device, sched_name):
    """Driving the B200 objects exactly like the reference loop (pipeline :465-489) gives the fused result."""
    from stablediffusion_b200.models.stable_diffusion import SDModelWrapper
    from stablediffusion_b200.pipelines.sd_unified_pipeline import StableDiffusionUnifiedPipeline

    dev = cuda_device
    cfg, ours, _ = _make_unet(dev, **SMALL_UNET)
    vcfg, vae, _ = _make_vae(dev)
    model = SDModelWrapper(vae=vae, base=ours, device=str(dev))
    model.set_scheduler(sched_name)
    stochastic = model.scheduler._stochastic
    g = torch.Generator().manual_seed(4)
    lat = torch.randn(1, 4, 16, 16, generator=g).to(dev)
    pe = torch.randn(1, 77, 64, generator=g).to(dev, torch.bfloat16)
    ne = torch.randn(1, 77, 64, generator=g).to(dev, torch.bfloat16)
    pipe = StableDiffusionUnifiedPipeline(do_cfg=True, device=str(dev), output_type="latents")
    if stochastic:
        pipe._step_generator = torch.Generator(device=dev).manual_seed(5)
        draws = torch.randn((6, 1, 4, 256), generator=torch.Generator(device=dev).manual_seed(5), device=dev)

        class _Replay:  # hands scheduler.step the draw the fused engine used for the same step
            def __init__(self):
                self.i = 0
        replay = _Replay()
        real_randn = torch.randn

        def fake_randn(shape, generator=None, **kw):
            if generator is replay:
                replay.i += 1
                return draws[replay.i - 1].view(shape).to(kw.get("dtype", torch.float32))
            return real_randn(shape, generator=generator, **kw)
    fused = pipe(model, prompt_embeds=pe, negative_prompt_embeds=ne, num_inference_steps=6, guidance_scale=5.0,
                 latents=lat, height=128, width=128)
    sch = model.scheduler
    sch.set_timesteps(6, device=dev)
    latents = lat.float() * float(sch.init_noise_sigma)
    ctx = torch.cat([ne, pe])
    for t in sch.timesteps:
        inp = torch.cat([latents] * 2)
        inp = sch.scale_model_input(inp, t)
        noise_pred = model.base(inp, t, ctx, return_dict=False)[0]
        u, c = noise_pred.chunk(2)
        noise_pred = 5.0 * (c - u) + u
        if stochastic:
            torch.randn = fake_randn
            try:
                latents = sch.step(noise_pred, t, latents, return_dict=False, generator=replay)[0]
            finally:
                torch.randn = real_randn
        else:
            latents = sch.step(noise_pred, t, latents, return_dict=False)[0]
    rl2 = _rel_l2(latents.cpu(), fused.cpu())
    print(f"UNFUSED vs FUSED rel-L2 {rl2:.4g}")
    # the unfused pattern rounds eps and the scaled input to bf16 between the UNet, CFG and scheduler step, the fused loop
    # keeps them in fp32: 6 steps of a random-init UNet amplify that to ~2e-2 (the trajectory tolerance above is 5e-2)
    assert rl2 <= 3e-2


def test_pipeline_sd15_txt2img_20_euler(cuda_device):
    """BASELINE config 1 end to end (random-init SD1.5 + VAE): final image PSNR vs the fp32 oracle."""
    from oracle import pipeline as OP
    from oracle import schedulers as OS
    from stablediffusion_b200.models.stable_diffusion import SDModelWrapper
    from stablediffusion_b200.pipelines.sd_unified_pipeline import StableDiffusionUnifiedPipeline

    dev = cuda_device
    cfg, ours, oracle = _make_unet(dev)
    vcfg, vae, ovae = _make_vae(dev)
    model = SDModelWrapper(vae=vae, base=ours, device=str(dev))
    g = torch.Generator().manual_seed(1234)
    lat = torch.randn(1, 4, 64, 64, generator=g)
    pe = torch.randn(1, 77, 768, generator=g).to(torch.bfloat16)
    ne = torch.randn(1, 77, 768, generator=g).to(torch.bfloat16)
    pipe = StableDiffusionUnifiedPipeline(do_cfg=True, device=str(dev))
    img = pipe(model, prompt_embeds=pe.to(dev), negative_prompt_embeds=ne.to(dev), num_inference_steps=20,
               guidance_scale=7.5, latents=lat.to(dev)).float().cpu()
    assert img.shape == (1, 3, 512, 512)
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    osched = OS.make_scheduler("euler", dict(OS.SD15_SCHEDULER_CONFIG))
    osched.set_timesteps(20)
    x0 = (lat * float(osched.init_noise_sigma)).to(dev)
    zl = OP.denoise(oracle.to(dev), osched, x0, torch.cat([ne, pe]).float().to(dev), osched.timesteps, 7.5)
    with torch.no_grad():
        ref = ovae.to(dev).decode(zl / 0.18215).float().cpu()
    lat_ours = StableDiffusionUnifiedPipeline(do_cfg=True, device=str(dev), output_type="latents")(
        model, prompt_embeds=pe.to(dev), negative_prompt_embeds=ne.to(dev), num_inference_steps=20, guidance_scale=7.5,
        latents=lat.to(dev)).float().cpu()
    psnr = _psnr(img, ref, peak=(ref.max() - ref.min()).item())
    print(f"E2E SD1.5 20 Euler: latent rel-L2 {_rel_l2(lat_ours, zl.cpu()):.4g}, image PSNR {psnr:.1f} dB, launches/step {pipe.last_engine.launches_per_step}")
    assert torch.isfinite(img).all() and psnr >= 30.0


@pytest.mark.parametrize("sched_name,strength", [("euler", 1.0), ("DDIM", 0.6), ("DPM++ 2M", 0.75), ("PNDM", 1.0), ("euler_a", 0.5), ("uni_pc", 0.7)])
def test_inpaint_loop_vs_oracle(cuda_device, sched_name, strength):
    """Third pipeline mode (reference :266-380, :492-506): 4-channel inpainting — the masked region is denoised, the rest
    is re-noised original latents after every step.  `image` is passed in latent space (skips the VAE so that both sides
    see the same image latents); the mask is given at pixel resolution and down-sampled (nearest) by the pipeline."""
    from oracle import pipeline as OP
    from oracle import schedulers as OS
    from stablediffusion_b200.models.stable_diffusion import SDModelWrapper
    from stablediffusion_b200.pipelines.sd_unified_pipeline import StableDiffusionUnifiedPipeline

    dev = cuda_device
    cfg, ours, oracle = _make_unet(dev, **SMALL_UNET)
    vcfg, vae, _ = _make_vae(dev)
    model = SDModelWrapper(vae=vae, base=ours, device=str(dev))
    model.set_scheduler(sched_name)
    g = torch.Generator().manual_seed(31)
    B, steps = 2, 10
    # (the pipeline casts the image latents to the prompt-embedding dtype, reference :872: use bf16-exact values)
    image_latents = torch.randn(B, 4, 16, 16, generator=g).bfloat16().float()
    noise = torch.randn(B, 4, 16, 16, generator=g)
    pe = torch.randn(B, 77, 64, generator=g).to(torch.bfloat16)
    ne = torch.randn(B, 77, 64, generator=g).to(torch.bfloat16)
    mask_px = torch.zeros(1, 1, 128, 128)
    mask_px[:, :, 32:100, 20:90] = 1.0
    pipe = StableDiffusionUnifiedPipeline(do_cfg=True, device=str(dev), output_type="latents")
    step_noise = None
    osched = OS.make_scheduler(sched_name, dict(OS.SD15_SCHEDULER_CONFIG))
    osched.set_timesteps(steps)
    init_t = min(int(steps * strength), steps)
    ts = osched.timesteps[max(steps - init_t, 0) * osched.order:]
    if sched_name in OS.STOCHASTIC:
        pipe._step_generator = torch.Generator(device=dev).manual_seed(78)
        step_noise = torch.randn((len(ts), B, 4, 256), generator=torch.Generator(device=dev).manual_seed(78), device=dev,
                                 dtype=torch.float32).view(len(ts), B, 4, 16, 16).cpu()
    out = pipe(model, prompt_embeds=pe.to(dev), negative_prompt_embeds=ne.to(dev), num_inference_steps=steps,
               guidance_scale=6.0, latents=noise.to(dev), image=image_latents.to(dev), mask_image=mask_px, strength=strength,
               height=128, width=128).float().cpu()
    assert pipe.is_inpaint
    # oracle: reference :848-913 with `latents` given -> noise = latents, start = noise * init_noise_sigma
    x0 = noise * float(osched.init_noise_sigma)
    m = torch.nn.functional.interpolate(mask_px, size=(16, 16)).repeat(B, 1, 1, 1)
    ref = OP.denoise(oracle.cpu(), osched, x0, torch.cat([ne, pe]).float(), ts, 6.0, step_noise=step_noise,
                     inpaint=(m, image_latents, noise))
    rl2 = _rel_l2(out, ref)
    print(f"INPAINT PARITY {sched_name} strength {strength}: rel-L2 {rl2:.4g}")
    assert torch.isfinite(out).all() and rl2 <= 3.3e-2   # measured <= 2.2e-2 (round 1): 1.5x head-room
    # outside the mask the result IS the original image latents (last step: no re-noising)
    keep = (m == 0).expand_as(out)
    torch.testing.assert_close(out[keep], image_latents[keep], rtol=1e-6, atol=1e-6)
