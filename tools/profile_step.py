"""Per-op device times of one fused sampler step (CUDA events around every op; profiling aid, not a bench)."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from stablediffusion_b200 import _native as N
from stablediffusion_b200.models.stable_diffusion import SDModelWrapper
from stablediffusion_b200.pipelines.sd_unified_pipeline import StableDiffusionUnifiedPipeline

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="sd15"); ap.add_argument("--batch", type=int, default=8); ap.add_argument("--res", type=int, default=512)
ap.add_argument("--out", default="gpurun_out/profile_step.txt"); ap.add_argument("--top", type=int, default=400)
a = ap.parse_args()
dev = "cuda:0"
m = SDModelWrapper.from_random(a.model, device=dev)
pipe = StableDiffusionUnifiedPipeline(device=dev, output_type="latents")
D = 2048 if a.model == "sdxl" else 768
kw = dict(prompt_embeds=torch.randn(a.batch, 77, D, device=dev).bfloat16(), negative_prompt_embeds=torch.randn(a.batch, 77, D, device=dev).bfloat16(),
          num_inference_steps=3, height=a.res, width=a.res, seed=0)
if a.model == "sdxl":
    kw.update(pooled_prompt_embeds=torch.randn(a.batch, 1280, device=dev).bfloat16(), negative_pooled_prompt_embeds=torch.randn(a.batch, 1280, device=dev).bfloat16())
pipe(m, **kw)
eng = pipe.last_engine
best = None
for _ in range(5):
    ms = eng.run_unet_timed()
    best = ms if best is None else [min(x, y) for x, y in zip(best, ms)]
kinds = eng.b.prog.kinds
# whole-step time as the pipeline runs it (one CUDA-graph launch per step), to compare with the sum of per-op times
st = eng._stream
with torch.cuda.stream(st):
    for _ in range(3): eng.step_once(st)
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record(st)
    for _ in range(10): eng.step_once(st)
    g1.record(st)
torch.cuda.synchronize()
graph_ms = g0.elapsed_time(g1) / 10
lines, agg = [], {}
for i, (t, k) in enumerate(zip(best, kinds)):
    md = eng.b.meta.get(i, {})
    name = md.get("op", N.OP_NAMES[k])
    tf = md["flops"] / (t * 1e-3) / 1e12 if "flops" in md and t > 0 else None
    gbs = md["bytes"] / (t * 1e-3) / 1e9 if "bytes" in md and t > 0 else None
    key = name + (":" + "x".join(str(md.get(q)) for q in ("M", "N", "K")) if "M" in md else "") + (f":{md.get('Nq')}x{md.get('Nkv')}d{md.get('d')}" if name == "attention" else "") + (f":{md.get('HW', md.get('rows'))}x{md.get('C')}" if "bytes" in md else "")
    e = agg.setdefault(key, [0, 0.0, 0.0, 0.0])
    e[0] += 1; e[1] += t; e[2] += md.get("flops", 0); e[3] += md.get("bytes", 0)
tot = sum(best)
out = [f"# {a.model} res {a.res} batch {a.batch} (UNet batch {2*a.batch}); total {tot:.3f} ms over {len(best)} ops; min of 5 event-timed runs; graph replay {graph_ms:.3f} ms/step"]
for key, (n, t, fl, by) in sorted(agg.items(), key=lambda x: -x[1][1])[: a.top]:
    extra = f"{fl / (t * 1e-3) / 1e12:8.1f} TFLOP/s" if fl else (f"{by / (t * 1e-3) / 1e9:8.1f} GB/s" if by else "")
    out.append(f"{t:8.3f} ms {100 * t / tot:5.1f}% n={n:3d} {t / n * 1000:8.1f} us/op {extra}  {key}")
open(a.out, "w").write("\n".join(out) + "\n")
print("\n".join(out[:60]))
