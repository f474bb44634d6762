"""Where does one 8-image job (SD1.5 512^2, DDIM 50, CFG) spend its time?  Whole pipeline call vs the 50-step loop vs the VAE
decode, device-timed (tuning aid)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import contextlib
import torch
import bench
from stablediffusion_b200.models.stable_diffusion import SDModelWrapper
from stablediffusion_b200.pipelines.sd_unified_pipeline import StableDiffusionUnifiedPipeline

dev = torch.device("cuda:0")
model_type, H, W, B, sampler, steps, guidance, _ = bench.WORKLOADS["sd15_512_b8_ddim50_cfg"]
model = SDModelWrapper.from_random(model_type, seed=0, device=str(dev))
with contextlib.redirect_stdout(sys.stderr):
    model.set_scheduler(sampler)
pipe = StableDiffusionUnifiedPipeline(do_cfg=True, device=str(dev))
inp = {k: v.to(dev) for k, v in bench.synth_inputs(model_type, B, H, W, seed=1).items()}
kw = dict(prompt_embeds=inp["pe"], negative_prompt_embeds=inp["ne"], latents=inp["lat"], num_inference_steps=steps,
          guidance_scale=guidance, height=H, width=W)
def ev():
    return torch.cuda.Event(enable_timing=True)
for _ in range(2):
    img = pipe(model, **kw)
torch.cuda.synchronize()
e0, e1 = ev(), ev()
t0 = time.perf_counter(); e0.record(); img = pipe(model, **kw); e1.record(); t_host = time.perf_counter() - t0
torch.cuda.synchronize()
print(f"whole call: {e0.elapsed_time(e1):.1f} ms on the device, host returned after {t_host*1e3:.1f} ms")
eng = pipe.last_engine
e0, e1 = ev(), ev()
e0.record(); eng.counter.zero_(); eng.run(steps); e1.record(); torch.cuda.synchronize()
print(f"{steps}-step loop (prep + {steps} graph launches): {e0.elapsed_time(e1):.1f} ms = {e0.elapsed_time(e1)/steps:.3f} ms/step")
lat = eng.latents.clone()
e0, e1 = ev(), ev()
e0.record(); out = model.vae.decode(lat / model.vae.config.scaling_factor, return_dict=False)[0]; e1.record(); torch.cuda.synchronize()
print(f"vae.decode call: {e0.elapsed_time(e1):.1f} ms")
st = [v for k, v in model.vae._programs.items() if k[0] == "dec"][0]
s = torch.cuda.current_stream().cuda_stream
e0, e1 = ev(), ev()
e0.record(); st["b"].prog.run(s); e1.record(); torch.cuda.synchronize()
print(f"vae decode program alone: {e0.elapsed_time(e1):.1f} ms over {st['b'].prog.num_launches} launches")
