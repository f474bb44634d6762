"""2-GPU check of the multi-GPU paths (launch under torch.distributed.run, one rank per GPU):
  * cond/uncond pair split (BASELINE config 5): both ranks of a pair must reproduce the latents of the
    unsplit single-GPU loop (same kernels, same order of fp32 operations in CFG + scheduler step; the UNet runs at
    half the batch, which may change GEMM tilings, hence a small tolerance),
  * data parallel sharding: rank r's images == rows of the full-batch run,
and a timing of the split vs the unsplit step loop (device events, max over ranks)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from stablediffusion_b200.distributed import make_pair_group, shard_range
from stablediffusion_b200.models.stable_diffusion import SDModelWrapper
from stablediffusion_b200.pipelines.sd_unified_pipeline import StableDiffusionUnifiedPipeline

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
B, steps = int(os.environ.get("CHECK_B", "2")), int(os.environ.get("CHECK_STEPS", "10"))
model = SDModelWrapper.from_random("sd15", seed=0, device=str(dev))
g = torch.Generator().manual_seed(7)
pe = torch.randn(B, 77, 768, generator=g).to(torch.bfloat16).to(dev)
ne = torch.randn(B, 77, 768, generator=g).to(torch.bfloat16).to(dev)
lat = torch.randn(B, 4, 64, 64, generator=g).to(dev)
kw = dict(num_inference_steps=steps, guidance_scale=7.5, height=512, width=512)

def timed(pipe, **k):
    pipe(model, **k)  # warm-up (graph capture)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); out = pipe(model, **k); e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return out, t.item()

ref_pipe = StableDiffusionUnifiedPipeline(do_cfg=True, device=str(dev), output_type="latents")
ref, t_ref = timed(ref_pipe, prompt_embeds=pe, negative_prompt_embeds=ne, latents=lat, **kw)
grp, pr = make_pair_group()
split_pipe = StableDiffusionUnifiedPipeline(do_cfg=True, device=str(dev), output_type="latents", pair_group=grp, pair_rank=pr)
got, t_split = timed(split_pipe, prompt_embeds=pe, negative_prompt_embeds=ne, latents=lat, **kw)
rel = ((got.float() - ref.float()).norm() / ref.float().norm()).item()
# data parallel: this rank's shard of the batch
lo, hi = shard_range(B, rank, world)
dp, _ = timed(ref_pipe, prompt_embeds=pe[lo:hi], negative_prompt_embeds=ne[lo:hi], latents=lat[lo:hi], **kw) if hi > lo else (ref[:0], 0.0)
rel_dp = ((dp.float() - ref[lo:hi].float()).norm() / (ref[lo:hi].float().norm() + 1e-12)).item() if hi > lo else 0.0
res = torch.tensor([rel, rel_dp], device=dev); dist.all_reduce(res, op=dist.ReduceOp.MAX)
if rank == 0:
    print(json.dumps({"world": world, "B": B, "steps": steps, "pair_split_rel_l2_vs_unsplit": res[0].item(), "dp_shard_rel_l2_vs_full_batch": res[1].item(),
                      "loop_ms_unsplit_1gpu": t_ref, "loop_ms_pair_split_2gpu": t_split}))
    assert res[0].item() < 5e-2 and res[1].item() < 5e-2  # the trajectory tolerance of tests/test_models_gpu.py (batch size changes GEMM tilings)
dist.destroy_process_group()
