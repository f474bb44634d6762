"""Multi-GPU equivalence check (launch under torch.distributed.run, one rank per GPU, world size 2 or more, even):
  * data parallel: rank r's shard of the batch vs rows [lo, hi) of the full-batch run on one GPU
  * cond/uncond pair split (BASELINE config 5): ranks (2p, 2p+1) run one half of the CFG batch each and exchange the eps
    halves once per step (engine.DenoiseEngine); result vs the unsplit single-GPU loop
Both in the default mode (tilings follow the batch: close, tolerance stated) and in the batch-invariant mode
(b200sd_set_batch_invariant: reduction orders follow the per-image shape — both must be BIT-IDENTICAL).
Also times the unsplit and the pair-split loop (device events, max over ranks).  Rank 0 prints one JSON line.
tests/test_multigpu_gpu.py runs this and asserts on the line."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from stablediffusion_b200 import _native as N
from stablediffusion_b200.distributed import make_pair_group, shard_range
from stablediffusion_b200.models.stable_diffusion import SDModelWrapper
from stablediffusion_b200.pipelines.sd_unified_pipeline import StableDiffusionUnifiedPipeline

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
B = int(os.environ.get("CHECK_B", "4"))
steps = int(os.environ.get("CHECK_STEPS", "10"))
res_px = int(os.environ.get("CHECK_RES", "512"))
exchange = os.environ.get("CHECK_EXCHANGE", "auto")
model = SDModelWrapper.from_random("sd15", seed=0, device=str(dev))
model.set_scheduler("DDIM")
g = torch.Generator().manual_seed(7)
pe = torch.randn(B, 77, 768, generator=g).to(torch.bfloat16).to(dev)
ne = torch.randn(B, 77, 768, generator=g).to(torch.bfloat16).to(dev)
lat = torch.randn(B, 4, res_px // 8, res_px // 8, generator=g).to(dev)
kw = dict(num_inference_steps=steps, guidance_scale=7.5, height=res_px, width=res_px)
grp, pr = make_pair_group()


def timed(pipe, **k):
    pipe(model, **k)  # warm-up (graph capture)
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = pipe(model, **k)
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return out.float(), t.item()


def rel(a, b):
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


out = {"world": world, "B": B, "steps": steps, "resolution": res_px}
for invariant in (False, True):
    N.set_batch_invariant(invariant)
    tag = "invariant" if invariant else "default"
    ref_pipe = StableDiffusionUnifiedPipeline(do_cfg=True, device=str(dev), output_type="latents")
    ref, t_ref = timed(ref_pipe, prompt_embeds=pe, negative_prompt_embeds=ne, latents=lat, **kw)
    split_pipe = StableDiffusionUnifiedPipeline(do_cfg=True, device=str(dev), output_type="latents", pair_group=grp, pair_rank=pr,
                                                pair_exchange=exchange)
    got, t_split = timed(split_pipe, prompt_embeds=pe, negative_prompt_embeds=ne, latents=lat, **kw)
    lo, hi = shard_range(B, rank, world)
    dp = timed(ref_pipe, prompt_embeds=pe[lo:hi], negative_prompt_embeds=ne[lo:hi], latents=lat[lo:hi], **kw)[0] if hi > lo else ref[:0]
    v = torch.tensor([rel(got, ref), rel(dp, ref[lo:hi]) if hi > lo else 0.0,
                      0.0 if torch.equal(got, ref) else 1.0, 0.0 if torch.equal(dp, ref[lo:hi]) else 1.0], device=dev)
    dist.all_reduce(v, op=dist.ReduceOp.MAX)
    out[tag] = {"pair_split_rel_l2_vs_unsplit": v[0].item(), "dp_shard_rel_l2_vs_full_batch": v[1].item(),
                "pair_split_bit_identical": v[2].item() == 0.0, "dp_shard_bit_identical": v[3].item() == 0.0,
                "loop_ms_unsplit_1gpu": t_ref, "loop_ms_pair_split": t_split,
                "pair_exchange": split_pipe.last_engine.exchange_kind,
                "exchange_bytes_per_step": B * 4 * (res_px // 8) ** 2 * 2}
N.set_batch_invariant(False)
if rank == 0:
    print(json.dumps(out))
dist.destroy_process_group()
