#!/usr/bin/env python
"""Benchmark of the denoising hot path (BASELINE.json metric: SD1.5 512^2 UNet steps/s and images/s).

One bench "step" = one full txt2img job of the workload's batch on every rank: `sampler_steps` fused
UNet+CFG+scheduler steps followed by the VAE decode (config[1] of BASELINE.json: SD1.5 512x512, batch 8 per GPU
with CFG, DDIM 50 steps; data-parallel over the GPUs = independent prompt shards, no collective in the loop).
  value            images/s, whole job over all ranks, inputs resident in HBM (device-event timed, max over ranks)
  e2e              the same jobs through the public pipeline call with HOST (pinned) inputs and a host image back
  roofline         tcgen05 GEMM kernel: algorithmic FLOPs of its launches / their CUDA-event time, vs MEASURED_PEAKS
  cpu_baseline     the oracle (fp32 restatement of the reference's diffusers path) on the host cores, bounded sample
`--impl reference` times that CPU path alone (the reference has no other implementation of this path: it is
100 % Python over diffusers, which is not installable here — see DESIGN.md).
Synthetic prompts (random embeddings) and random-init weights of the named architecture: no network here.
"""
import argparse
import contextlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# stdout carries exactly ONE JSON line.  Native libraries write to file descriptor 1 behind Python's back (NCCL prints its
# "NCCL version ..." banner with printf when NCCL_DEBUG is set in the environment, as it is on the GPU boxes; NCCL_DEBUG_FILE
# does not catch it), so main() points descriptor 1 at stderr for the whole run and the result line is written to the saved
# original descriptor.
_RESULT_FD = None


def capture_stdout() -> None:
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)


def emit_result(line: dict) -> None:
    sys.stdout.flush()
    if _RESULT_FD is None:
        print(json.dumps(line), flush=True)
    else:
        os.write(_RESULT_FD, (json.dumps(line) + "\n").encode())

WORKLOADS = {
    # name: (model_type, H, W, batch per GPU, scheduler, sampler steps, guidance)
    "sd15_512_b8_ddim50_cfg": ("sd15", 512, 512, 8, "DDIM", 50, 7.5),
    "sd15_512_b1_euler20_cfg": ("sd15", 512, 512, 1, "euler", 20, 7.5),
    "sdxl_1024_b4_euler30_cfg": ("sdxl", 1024, 1024, 4, "euler", 30, 5.0),
}
UNET_GFLOP = {("sd15", 512): 803.3, ("sd15", 768): 2148.1, ("sd15", 1024): 4674.0, ("sdxl", 1024): 6761.2, ("sdxl", 512): 1588.8}
VAE_GFLOP = {512: 2514.5, 768: 5754.3, 1024: 10470.4}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf=d.get("bf16_tflops_sustained", d["bf16_tflops"]), src="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf=1400.0, src="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        self.rows, self.stop, self.index = [], threading.Event(), index
        self.t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self.stop.wait(0.2)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.t.join(timeout=6)

    def summary(self):
        sm = sorted(float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit())
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 3 + i and r[3 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.rows)}


def synth_inputs(model_type, B, H, W, seed=1):
    import torch

    g = torch.Generator().manual_seed(seed)
    D = 2048 if model_type == "sdxl" else 768
    d = dict(pe=torch.randn(B, 77, D, generator=g).to(torch.bfloat16), ne=torch.randn(B, 77, D, generator=g).to(torch.bfloat16),
             lat=torch.randn(B, 4, H // 8, W // 8, generator=torch.Generator().manual_seed(1234)))
    if model_type == "sdxl":
        d["pp"] = torch.randn(B, 1280, generator=g).to(torch.bfloat16)
        d["np"] = torch.randn(B, 1280, generator=g).to(torch.bfloat16)
    return d


def cpu_reference_sample(model_type, H, W, sampler, sampler_steps, guidance, threads=None):
    """The reference's CPU path (oracle restatement, fp32) on a BOUNDED sample: one CFG sampler step at batch 1
    (2 UNet sample-forwards + CFG + scheduler step) and one VAE decode, extrapolated to the full job."""
    import torch

    from oracle import pipeline as OP
    from oracle import schedulers as OS
    from oracle.unet import SD15_UNET, SDXL_UNET, UNet2DConditionOracle
    from oracle.vae import AutoencoderKLOracle
    from stablediffusion_b200.config import unet_config, vae_config
    from stablediffusion_b200.random_init import random_state_dict
    from stablediffusion_b200.unet import unet_param_shapes
    from stablediffusion_b200.vae import vae_param_shapes

    cores = threads or os.cpu_count() or 1
    torch.set_num_threads(cores)
    cache = cpu_reference_sample.__dict__.setdefault("models", {})
    if model_type not in cache:
        unet = UNet2DConditionOracle(**(SDXL_UNET if model_type == "sdxl" else SD15_UNET)).eval()
        unet.load_state_dict(random_state_dict(unet_param_shapes(unet_config(model_type)), seed=0), strict=True)
        vae = AutoencoderKLOracle(scaling_factor=vae_config(model_type)["scaling_factor"]).eval()
        vae.load_state_dict(random_state_dict(vae_param_shapes(vae_config(model_type)), seed=1), strict=True)
        cache[model_type] = (unet, vae)
    unet, vae = cache[model_type]
    inp = synth_inputs(model_type, 1, H, W)
    sched = OS.make_scheduler(sampler, dict(OS.SDXL_SCHEDULER_CONFIG if model_type == "sdxl" else OS.SD15_SCHEDULER_CONFIG))
    sched.set_timesteps(sampler_steps)
    added = None
    if model_type == "sdxl":
        added = dict(text_embeds=torch.cat([inp["np"], inp["pp"]]).float(),
                     time_ids=torch.tensor([[H, W, 0, 0, H, W]], dtype=torch.float32).repeat(2, 1))
    ctx = torch.cat([inp["ne"], inp["pe"]]).float()
    lat = inp["lat"] * float(sched.init_noise_sigma)
    t0 = time.perf_counter()
    lat = OP.denoise(unet, sched, lat, ctx, sched.timesteps[:1], guidance, added_cond_kwargs=added)
    t_step = time.perf_counter() - t0
    t0 = time.perf_counter()
    with torch.no_grad():
        OP.decode(vae, lat)
    t_vae = time.perf_counter() - t0
    sec_per_image = sampler_steps * t_step + t_vae
    return dict(value=1.0 / sec_per_image, unit="images/s", cores=cores, kind="port",
                sample=f"1 CFG sampler step at batch 1 ({t_step:.2f} s) + 1 VAE decode ({t_vae:.2f} s), fp32 oracle, "
                       f"extrapolated to {sampler_steps} steps",
                step_s=t_step, vae_s=t_vae)


def run_reference(args, wl, rank):
    model_type, H, W, B, sampler, sampler_steps, guidance = WORKLOADS[wl]
    if rank != 0:
        return
    vals = []
    for i in range(args.warmup + args.steps):
        r = cpu_reference_sample(model_type, H, W, sampler, sampler_steps, guidance)
        if i >= args.warmup:
            vals.append(r)
        if i == 0 and r["step_s"] * (args.warmup + args.steps) > 240:   # keep the whole run within minutes
            vals = [r]
            break
    v = sum(x["value"] for x in vals) / len(vals)
    cb = dict(vals[-1]); cb["value"] = v
    cb.pop("step_s"); cb.pop("vae_s")
    line = {"impl": "reference", "metric": "images/s", "value": v, "unit": "images/s", "n_gpus": args.gpus, "steps": len(vals),
            "warmup": args.warmup, "ms_per_step": 1000.0 / v, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic prompts (random embeddings), random-init weights",
            "config": {"workload": wl, "note": "CPU oracle of the reference's diffusers path; per-image rate (batch 1)"},
            "cpu_baseline": cb, "e2e": {"value": v, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit_result(line)


def run_b200(args, wl, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    from stablediffusion_b200 import _native as N
    from stablediffusion_b200.models.stable_diffusion import SDModelWrapper
    from stablediffusion_b200.pipelines.sd_unified_pipeline import StableDiffusionUnifiedPipeline

    model_type, H, W, B, sampler, sampler_steps, guidance = WORKLOADS[wl]
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a B200: stablediffusion_b200 has no CPU fallback (use --impl reference for the CPU path)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    model = SDModelWrapper.from_random(model_type, seed=0, device=str(dev))
    with contextlib.redirect_stdout(sys.stderr):   # the wrapper mirrors the reference's print(); stdout carries ONE JSON line
        model.set_scheduler(sampler)
    pipe = StableDiffusionUnifiedPipeline(do_cfg=True, device=str(dev))
    # every rank owns an independent prompt shard (weak scaling: B images per GPU); inputs differ per rank
    inp = synth_inputs(model_type, B, H, W, seed=1 + rank)
    host = {k: v.pin_memory() for k, v in inp.items()}
    devin = {k: v.to(dev) for k, v in inp.items()}

    def job(src, to_host):
        kw = dict(prompt_embeds=src["pe"].to(dev, non_blocking=True), negative_prompt_embeds=src["ne"].to(dev, non_blocking=True),
                  latents=src["lat"].to(dev, non_blocking=True), num_inference_steps=sampler_steps, guidance_scale=guidance,
                  height=H, width=W)
        if model_type == "sdxl":
            kw.update(pooled_prompt_embeds=src["pp"].to(dev, non_blocking=True), negative_pooled_prompt_embeds=src["np"].to(dev, non_blocking=True))
        img = pipe(model, **kw)
        if to_host:
            himg = img.to("cpu", non_blocking=False)
            return himg
        return img

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(src, to_host, iters):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(iters):
            out = job(src, to_host)
        e1.record()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        dev_s = e0.elapsed_time(e1) / 1000.0
        t = torch.tensor([dev_s, wall], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        barrier()
        return t[0].item(), t[1].item(), out

    for _ in range(max(args.warmup, 1)):
        job(devin, False)
    with ClockSampler(local_rank) as clk:
        dev_s, _, _ = timed(devin, False, args.steps)
    for _ in range(1):
        job(host, True)
    _, e2e_wall, himg = timed(host, True, args.steps)

    eng = pipe.last_engine
    vae_prog = [v for k, v in model.vae._programs.items() if k[0] == "dec"][0]["b"].prog
    launches_per_job = 1 + sampler_steps * eng.launches_per_step + vae_prog.num_launches
    images = B * world * args.steps
    value = images / dev_s
    e2e_value = images / e2e_wall

    # ---- roofline of the dominant kernel (tcgen05 GEMM): per-op CUDA events over one step program -------------
    pk = peaks()
    roof = None
    if rank == 0:
        ms = eng.run_unet_timed()
        ms = eng.run_unet_timed()
        kinds = eng.b.prog.kinds
        gemm_ms = sum(m for m, k in zip(ms, kinds) if k == N.OP_GEMM)
        att_ms = sum(m for m, k in zip(ms, kinds) if k == N.OP_ATTENTION)
        n_gemm = sum(1 for k in kinds if k == N.OP_GEMM)
        gemm_flops = eng.gemm_flops
        achieved = gemm_flops / (gemm_ms * 1e-3) / 1e12
        traffic, traffic_src = None, None   # DRAM bytes per launch of the same kernel from the committed ncu capture
        tp = os.path.join(ROOT, "profiles", "r01_gemm_traffic.json")
        if wl == "sd15_512_b8_ddim50_cfg" and os.path.exists(tp):
            with open(tp) as f:
                td = json.load(f)
            traffic, traffic_src = td["dram_bytes_per_launch"], "profiles/r01_gemm_traffic.json (ncu, cold cache per launch)"
        roof = {"bound": "tensor", "kernel": "gemm_tc_kernel (tcgen05 implicit GEMM, all conv/linear launches of one sampler step)",
                "achieved": achieved, "peak": pk["tf"], "unit": "TFLOP/s", "frac": achieved / pk["tf"], "traffic": traffic, "traffic_source": traffic_src,
                "algorithmic_flop_per_launch": gemm_flops / n_gemm,
                "peak_source": f"{pk['src']} bf16_tflops_sustained", "launches": n_gemm, "avg_launch_ms": gemm_ms / n_gemm,
                "step_ms_by_kind": {N.OP_NAMES[k]: round(sum(m for m, kk in zip(ms, kinds) if kk == k), 4) for k in sorted(set(kinds))},
                "attention_tflops": (eng.attn_flops / (att_ms * 1e-3) / 1e12) if att_ms > 0 else None}
    unet_gf = UNET_GFLOP.get((model_type, H))
    job_tflop = (sampler_steps * 2 * B * unet_gf + B * VAE_GFLOP[H]) / 1000.0 if unet_gf else None
    if rank == 0:
        cb = None
        if world == 1 and not args.no_cpu_baseline:
            cb = cpu_reference_sample(model_type, H, W, sampler, sampler_steps, guidance)
            cb.pop("step_s"); cb.pop("vae_s")
        h2d = sum(v.numel() * v.element_size() for v in host.values())
        line = {"metric": "images/s", "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": 1000.0 * dev_s / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "bf16", "data": "synthetic prompts (random embeddings), random-init weights of the named architecture",
                "config": {"workload": wl, "model": model_type, "resolution": f"{H}x{W}", "batch_per_gpu": B, "scheduler": sampler,
                           "sampler_steps": sampler_steps, "guidance_scale": guidance, "cfg": True, "parallelism": f"dp{world}",
                           "l2": "working set (1.7 GB weights + activations) exceeds the 126 MB L2; no explicit flush"},
                "unet_steps_per_s": sampler_steps * world * args.steps / dev_s,
                "algorithmic_tflop_per_job": job_tflop,
                "tensor_frac_whole_job": (job_tflop * args.steps / dev_s / pk["tf"]) if job_tflop else None,
                "e2e": {"value": e2e_value, "unit": "images/s", "h2d_bytes_per_step": h2d,
                        "d2h_bytes_per_step": himg.numel() * himg.element_size()},
                "gpu_launches": launches_per_job * args.steps, "launches_per_sampler_step": eng.launches_per_step,
                "clocks": clk.summary(), "roofline": roof, "cpu_baseline": cb}
        emit_result(line)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="sd15_512_b8_ddim50_cfg", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    capture_stdout()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, args.workload, rank)
        return
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run --nproc-per-node {args.gpus}")
    run_b200(args, args.workload, rank, world, local_rank)


if __name__ == "__main__":
    main()
