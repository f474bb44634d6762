"""BASELINE config 5: UNet-only and VAE-only throughput sweep, batch 1-64 at 512^2 and 1024^2 (SD1.5), on one GPU and
— launched under torch.distributed.run with 2 ranks — with the cond/uncond GPU-pair split.

    python tools/sweep_config5.py                                  # one GPU
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/sweep_config5.py   # + pair split

Per (batch, resolution): ms per CFG sampler step of the fused loop (one CUDA-graph launch: UNet at batch 2B + CFG +
scheduler update; with the split: UNet at batch B per GPU + the in-graph exchange), UNet steps/s, its algorithmic tensor
fraction, and ms per VAE decode of B images.  Shapes that do not fit the GPU's memory are reported as "oom" (the CUDA
out-of-memory exception is caught; nothing is forced).  Prints a table and one JSON line."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from bench import UNET_GFLOP, peaks
from stablediffusion_b200.engine import DenoiseEngine
from stablediffusion_b200.models.stable_diffusion import SDModelWrapper

rank, world, lr = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
pair = None
if world > 1:
    from stablediffusion_b200.distributed import make_pair_group

    dist.init_process_group("nccl", device_id=dev)
    pair = make_pair_group()
batches = [int(x) for x in os.environ.get("SWEEP_B", "1,2,4,8,16,32,64").split(",")]
resolutions = [int(x) for x in os.environ.get("SWEEP_RES", "512,1024").split(",")]
steps = int(os.environ.get("SWEEP_STEPS", "8"))
model = SDModelWrapper.from_random("sd15", seed=0, device=str(dev))
model.set_scheduler("DDIM")
pk = peaks()
rows = []


def unet_step_ms(B, h, split):
    kw = dict(pair_group=pair[0], pair_rank=pair[1], pair_exchange=os.environ.get("SWEEP_EXCHANGE", "auto")) if split else {}
    eng = DenoiseEngine(model.base, B, h, h, True, 768, False, **kw)
    sch = model.scheduler
    sch.set_timesteps(50, device=dev)
    g = torch.Generator().manual_seed(3)
    lat = torch.randn(B, 4, h, h, generator=g).to(dev)
    ctx = torch.randn(2 * B, 77, 768, generator=g).to(torch.bfloat16).to(dev)
    eng.load(lat, ctx, sch.rows_for(0, 50), 7.5)
    eng.run(3)                      # warm-up + graph capture
    best = 1e30
    for _ in range(3):
        if world > 1:
            dist.barrier()
        eng.run(steps)
        t = torch.tensor([eng.last_loop_ms() / steps], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        best = min(best, t.item())
    kind = eng.exchange_kind
    del eng
    torch.cuda.empty_cache()
    return best, kind


def vae_ms(B, h):
    z = torch.randn(B, 4, h, h, device=dev)
    for _ in range(2):
        model.vae.decode(z, scale=1 / 0.18215)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        model.vae.decode(z, scale=1 / 0.18215)
    e1.record()
    torch.cuda.synchronize()
    model.vae._programs.clear()
    torch.cuda.empty_cache()
    return e0.elapsed_time(e1) / 3


def guarded(fn, *a):
    try:
        return fn(*a)
    except torch.cuda.OutOfMemoryError:
        torch.cuda.empty_cache()
        return None
    except RuntimeError as e:
        if "out of memory" in str(e).lower():
            torch.cuda.empty_cache()
            return None
        raise


for res in resolutions:
    h = res // 8
    gf = UNET_GFLOP[("sd15", res)]
    for B in batches:
        r = dict(res=res, B=B)
        u = guarded(unet_step_ms, B, h, False)
        r["unet_step_ms"] = u[0] if u else None
        if u:
            r["unet_steps_per_s"] = 1000.0 / u[0]
            r["tensor_frac"] = 2 * B * gf / 1000.0 / (u[0] * 1e-3) / pk["tf"]
        if pair is not None:
            s = guarded(unet_step_ms, B, h, True)
            ok = torch.tensor([1 if s else 0], device=dev)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            r["pair_split_step_ms"] = s[0] if (s and ok.item()) else None
            r["pair_exchange"] = s[1] if s else None
            r["exchange_bytes_per_step"] = B * 4 * h * h * 2
        if rank == 0:
            v = guarded(vae_ms, B, h)
            r["vae_decode_ms"] = v
        rows.append(r)
        if rank == 0:
            print(json.dumps(r), file=sys.stderr, flush=True)
if rank == 0:
    print(f"# SD1.5 sweep (BASELINE config 5), {torch.cuda.get_device_name(0)}, world {world}; peak {pk['tf']} TFLOP/s ({pk['src']})")
    print(f"{'res':>5} {'B':>3} {'UNet step ms':>13} {'steps/s':>8} {'tensor frac':>11} {'pair-split ms':>13} {'speed-up':>8} {'VAE decode ms':>13}")
    for r in rows:
        f = lambda v, p=2: "oom" if v is None else f"{v:.{p}f}"
        sp = (r["unet_step_ms"] / r["pair_split_step_ms"]) if r.get("pair_split_step_ms") and r.get("unet_step_ms") else None
        print(f"{r['res']:>5} {r['B']:>3} {f(r.get('unet_step_ms')):>13} {f(r.get('unet_steps_per_s'), 1):>8} {f(r.get('tensor_frac'), 3):>11} "
              f"{f(r.get('pair_split_step_ms')) if pair is not None else '-':>13} {f(sp) if sp else '-':>8} {f(r.get('vae_decode_ms')):>13}")
    print(json.dumps({"sweep": rows, "world": world}))
if world > 1:
    dist.destroy_process_group()
