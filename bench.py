#!/usr/bin/env python
"""Benchmark of the denoising hot path (BASELINE.json metric: SD1.5 512^2 UNet steps/s and images/s at 1/2/4/8 B200;
% tensor-pipe / HBM roofline).

One bench "step" = one full txt2img job of the workload's batch on every rank: `sampler_steps` fused
UNet+CFG+scheduler steps (one CUDA-graph launch each) followed by the VAE decode.  Default workload = BASELINE.json
config[1]: SD1.5 512x512, batch 8 per GPU with CFG, DDIM 50 steps; data parallel over the GPUs = independent prompt shards,
no collective in the loop (weak scaling; `--strong` fixes the global batch instead, `--pair-split` runs the cond/uncond
GPU-pair split of config 5).
  value            images/s, whole job over all ranks, inputs resident in HBM (device-event timed, max over ranks)
  e2e              the same jobs through the public pipeline call with HOST (pinned) inputs and a host image back
  roofline         tcgen05 GEMM kernel: algorithmic FLOPs of its launches / their in-step time (the CUDA-graph step time
                   measured inside the timed region x the kernel's share of the step), vs MEASURED_PEAKS
  roofline_hbm     GroupNorm(+SiLU) and LayerNorm: algorithmic bytes (1R + 1W) / in-step time vs the measured HBM peak
  extra_workloads  configs 1 / 3 / 4 of BASELINE.json, a few jobs each (N = 1 only)
  gpu_yardstick    torch-eager bf16 of the same UNet step on the same GPU (cuDNN / cuBLASLt / SDPA) — a library number to
                   beat, NOT the reference
  cpu_baseline     the oracle (fp32 restatement of the reference's diffusers path) on the host cores, bounded sample
`--impl reference` times that CPU path alone (the reference has no other implementation of this path: it is 100 % Python
over diffusers, which is not installable here — see DESIGN.md): every timed "step" is a bounded sample of the workload
(a few CFG sampler steps at batch 1 + one VAE decode), BASELINE config 1 is additionally timed in full.
Synthetic prompts (random embeddings) and random-init weights of the named architecture: no network here.
"""
import argparse
import contextlib
import json
import os
import statistics
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
    # name: (model_type, H, W, batch per GPU, scheduler, scheduler steps, guidance, img2img strength or None)
    "sd15_512_b8_ddim50_cfg": ("sd15", 512, 512, 8, "DDIM", 50, 7.5, None),            # BASELINE config 2 (the headline)
    "sd15_512_b1_euler20_cfg": ("sd15", 512, 512, 1, "euler", 20, 7.5, None),          # config 1
    "sdxl_1024_b4_euler30_cfg": ("sdxl", 1024, 1024, 4, "euler", 30, 5.0, None),       # config 3
    "sd15_768_img2img_lora_s06": ("sd15", 768, 768, 1, "euler", 50, 7.5, 0.6),         # config 4: rank-16 LoRA merged, 30 of 50 steps
}
EXTRA_WORKLOADS = ("sd15_512_b1_euler20_cfg", "sdxl_1024_b4_euler30_cfg", "sd15_768_img2img_lora_s06")
UNET_GFLOP = {("sd15", 512): 803.3, ("sd15", 768): 2148.1, ("sd15", 1024): 4674.0, ("sdxl", 1024): 6761.2, ("sdxl", 512): 1588.8}
VAE_GFLOP = {512: 2514.5, 768: 5754.3, 1024: 10470.4}
VAE_ENC_GFLOP = {512: 1116.7, 768: 2609.1}


def executed_steps(steps, strength):
    return steps if strength is None else min(int(steps * strength), steps)


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


def synth_inputs(model_type, B, H, W, seed=1, image=False):
    import torch

    g = torch.Generator().manual_seed(seed)
    D = 2048 if model_type == "sdxl" else 768
    d = dict(pe=torch.randn(B, 77, D, generator=g).to(torch.bfloat16), ne=torch.randn(B, 77, D, generator=g).to(torch.bfloat16),
             lat=torch.randn(B, 4, H // 8, W // 8, generator=torch.Generator().manual_seed(1234)))
    if model_type == "sdxl":
        d["pp"] = torch.randn(B, 1280, generator=g).to(torch.bfloat16)
        d["np"] = torch.randn(B, 1280, generator=g).to(torch.bfloat16)
    if image:  # img2img input image U(-1, 1); the latents come from the VAE encoder then
        d["img"] = (torch.rand(B, 3, H, W, generator=torch.Generator().manual_seed(2)) * 2 - 1).to(torch.bfloat16)
        del d["lat"]
    return d


def rank16_lora(unet_cfg, seed=3, rank=16):
    """A LoRA file as pipelines/train_lora_pipeline.py:497-528 writes it (to_q/k/v/to_out.0 of every attention, alpha = r)."""
    import torch

    from stablediffusion_b200.unet import unet_param_shapes

    shapes = unet_param_shapes(unet_cfg)
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for k in shapes:
        if k.endswith((".to_q.weight", ".to_k.weight", ".to_v.weight", ".to_out.0.weight")):
            out_f, in_f = shapes[k]
            sd[f"unet.{k[:-7]}.lora.down.weight"] = torch.randn(rank, in_f, generator=g) / rank
            sd[f"unet.{k[:-7]}.lora.up.weight"] = torch.randn(out_f, rank, generator=g) * 0.02
    return sd


# ------------------------------------------------------------------------------------------------------
# CPU path (oracle port of the reference's diffusers calls): --impl reference and cpu_baseline
# ------------------------------------------------------------------------------------------------------
class CpuReference:
    def __init__(self, model_type, threads=None):
        import torch

        from oracle.unet import SD15_UNET, SDXL_UNET, UNet2DConditionOracle
        from oracle.vae import AutoencoderKLOracle
        from stablediffusion_b200.config import unet_config, vae_config
        from stablediffusion_b200.random_init import random_state_dict
        from stablediffusion_b200.unet import unet_param_shapes
        from stablediffusion_b200.vae import vae_param_shapes

        self.model_type = model_type
        self.cores = threads or os.cpu_count() or 1
        torch.set_num_threads(self.cores)
        self.unet = UNet2DConditionOracle(**(SDXL_UNET if model_type == "sdxl" else SD15_UNET)).eval()
        self.unet.load_state_dict(random_state_dict(unet_param_shapes(unet_config(model_type)), seed=0), strict=True)
        self.vae = AutoencoderKLOracle(scaling_factor=vae_config(model_type)["scaling_factor"]).eval()
        self.vae.load_state_dict(random_state_dict(vae_param_shapes(vae_config(model_type)), seed=1), strict=True)

    def _setup(self, H, W, sampler, sched_steps):
        import torch

        from oracle import schedulers as OS

        inp = synth_inputs(self.model_type, 1, H, W)
        sched = OS.make_scheduler(sampler, dict(OS.SDXL_SCHEDULER_CONFIG if self.model_type == "sdxl" else OS.SD15_SCHEDULER_CONFIG))
        sched.set_timesteps(sched_steps)
        added = None
        if self.model_type == "sdxl":
            added = dict(text_embeds=torch.cat([inp["np"], inp["pp"]]).float(),
                         time_ids=torch.tensor([[H, W, 0, 0, H, W]], dtype=torch.float32).repeat(2, 1))
        ctx = torch.cat([inp["ne"], inp["pe"]]).float()
        lat = inp["lat"] * float(sched.init_noise_sigma)
        return sched, ctx, lat, added

    def sample(self, H, W, sampler, sched_steps, guidance, n_sampler_steps, decode=True):
        """Wall time of `n_sampler_steps` CFG sampler steps at batch 1 (2 UNet sample-forwards + CFG + scheduler step each)
        and of one VAE decode."""
        import torch

        from oracle import pipeline as OP

        sched, ctx, lat, added = self._setup(H, W, sampler, sched_steps)
        t0 = time.perf_counter()
        lat = OP.denoise(self.unet, sched, lat, ctx, sched.timesteps[:n_sampler_steps], guidance, added_cond_kwargs=added)
        t_steps = time.perf_counter() - t0
        t_vae = 0.0
        if decode:
            t0 = time.perf_counter()
            with torch.no_grad():
                OP.decode(self.vae, lat)
            t_vae = time.perf_counter() - t0
        return t_steps, t_vae


def cpu_baseline_for(wl, budget_steps=8):
    """cpu_baseline of the GPU arm's line: a bounded sample of the SAME workload (about 10-20 s of CPU work)."""
    model_type, H, W, B, sampler, sched_steps, guidance, strength = WORKLOADS[wl]
    ref = CpuReference(model_type)
    n_exec = executed_steps(sched_steps, strength)
    n = min(budget_steps, n_exec)
    t_steps, t_vae = ref.sample(H, W, sampler, sched_steps, guidance, n)
    per_step = t_steps / n
    sec_per_image = n_exec * per_step + t_vae
    return dict(value=1.0 / sec_per_image, unit="images/s", cores=ref.cores, kind="port",
                sample=f"{n} CFG sampler steps at batch 1 ({t_steps:.2f} s) + 1 VAE decode ({t_vae:.2f} s) of {wl}, fp32 oracle on "
                       f"{ref.cores} host threads, extrapolated to {n_exec} steps per image")


def run_reference(args, wl, rank):
    """The reference arm: each of the K timed "steps" is a bounded sample of the workload — SAMPLE_STEPS CFG sampler steps at
    batch 1 plus one VAE decode — and `value` extrapolates one image from it (CPU throughput per image does not depend on the
    batch).  `ms_per_step` is the wall time of one sample, so steps x ms_per_step is what was really timed.  BASELINE config 1
    (batch 1, 20 Euler steps + decode) is also timed IN FULL, once, and reported under `config1_full`."""
    if rank != 0:
        return
    model_type, H, W, B, sampler, sched_steps, guidance, strength = WORKLOADS[wl]
    n_exec = executed_steps(sched_steps, strength)
    ref = CpuReference(model_type)
    sample_steps = 3
    budget_s = float(os.environ.get("B200SD_REF_BUDGET_S", "200"))
    t_begin = time.perf_counter()
    samples = []
    todo = args.warmup + args.steps
    i = 0
    while i < todo:
        t0 = time.perf_counter()
        t_steps, t_vae = ref.sample(H, W, sampler, sched_steps, guidance, sample_steps)
        wall = time.perf_counter() - t0
        if i == 0:  # size the run to the box: fewer sampler steps per sample, then fewer warm-ups, then fewer samples
            if wall * todo > budget_s and sample_steps > 1:
                sample_steps = 1
            est = wall * (sample_steps / 3.0 + 0.3)
            if est * todo > budget_s:
                todo = max(2, int(budget_s / est))
        if i >= min(args.warmup, todo - 1):
            samples.append(dict(t_steps=t_steps, t_vae=t_vae, wall=wall, n=sample_steps if i else 3))
        i += 1
    per_image = [n_exec * s["t_steps"] / s["n"] + s["t_vae"] for s in samples]
    vals = [1.0 / p for p in per_image]
    v = statistics.mean(vals)
    line = {"impl": "reference", "metric": "images/s", "value": v, "unit": "images/s", "n_gpus": args.gpus, "steps": len(samples),
            "warmup": min(args.warmup, todo - 1), "ms_per_step": 1000.0 * statistics.mean(s["wall"] for s in samples),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic prompts (random embeddings), random-init weights of the named architecture",
            "config": {"workload": wl, "model": model_type, "resolution": f"{H}x{W}", "scheduler": sampler, "sampler_steps": n_exec,
                       "guidance_scale": guidance, "cfg": True,
                       "note": "CPU oracle port of the reference's diffusers path; per-image rate at batch 1"},
            "spread": {"min": min(vals), "max": max(vals), "stdev": statistics.pstdev(vals), "samples": len(vals)},
            "cpu_baseline": {"value": v, "unit": "images/s", "cores": ref.cores, "kind": "port",
                             "sample": f"each timed step = {samples[-1]['n']} CFG sampler steps at batch 1 + 1 VAE decode of {wl} "
                                       f"(fp32 oracle, {ref.cores} host threads), extrapolated to {n_exec} sampler steps per image"},
            "e2e": {"value": v, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    # BASELINE config 1 in full: 20 Euler steps at batch 1 + the decode, nothing extrapolated
    if not args.no_config1 and time.perf_counter() - t_begin < budget_s + 60:
        m1, H1, W1, _, s1, n1, g1, _ = WORKLOADS["sd15_512_b1_euler20_cfg"]
        ref1 = ref if m1 == model_type else CpuReference(m1)
        t_steps, t_vae = ref1.sample(H1, W1, s1, n1, g1, n1)
        line["config1_full"] = {"workload": "sd15_512_b1_euler20_cfg", "images_per_s": 1.0 / (t_steps + t_vae), "seconds": t_steps + t_vae,
                                "sampler_steps_timed": n1, "unet_s": t_steps, "vae_decode_s": t_vae, "extrapolated": False}
    emit_result(line)


# ------------------------------------------------------------------------------------------------------
# the B200 arm
# ------------------------------------------------------------------------------------------------------
def build_model(wl, dev):
    from stablediffusion_b200.models.stable_diffusion import SDModelWrapper

    model_type, H, W, B, sampler, sched_steps, guidance, strength = WORKLOADS[wl]
    model = SDModelWrapper.from_random(model_type, seed=0, device=str(dev), with_vae_encoder=strength is not None)
    with contextlib.redirect_stdout(sys.stderr):   # the wrapper mirrors the reference's print(); stdout carries ONE JSON line
        model.set_scheduler(sampler)
        if "lora" in wl:  # config 4: a rank-16 adapter in train_lora_pipeline's format, merged on the device
            model.load_lora_weights(rank16_lora(model.base.config), adapter_name="bench_r16")
            model.set_adapters(["bench_r16"], [1.0])
    return model


def make_job(wl, model, pipe, dev, B):
    model_type, H, W, _, sampler, sched_steps, guidance, strength = WORKLOADS[wl]

    def job(src, to_host):
        kw = dict(prompt_embeds=src["pe"].to(dev, non_blocking=True), negative_prompt_embeds=src["ne"].to(dev, non_blocking=True),
                  num_inference_steps=sched_steps, guidance_scale=guidance, height=H, width=W)
        if strength is None:
            kw["latents"] = src["lat"].to(dev, non_blocking=True)
        else:
            kw.update(image=src["img"].to(dev, non_blocking=True), strength=strength, seed=1234)
        if model_type == "sdxl":
            kw.update(pooled_prompt_embeds=src["pp"].to(dev, non_blocking=True), negative_pooled_prompt_embeds=src["np"].to(dev, non_blocking=True))
        img = pipe(model, **kw)
        return img.to("cpu", non_blocking=False) if to_host else img

    return job


def job_tflop(wl, B):
    model_type, H, W, _, sampler, sched_steps, guidance, strength = WORKLOADS[wl]
    unet_gf = UNET_GFLOP.get((model_type, H))
    if unet_gf is None:
        return None
    n_exec = executed_steps(sched_steps, strength)
    enc = VAE_ENC_GFLOP.get(H, 0.0) if strength is not None else 0.0
    return (n_exec * 2 * B * unet_gf + B * (VAE_GFLOP[H] + enc)) / 1000.0


def step_breakdown(eng, N, passes=5):
    """Median per-kind share of one sampler step from `passes` per-op CUDA-event passes (each preceded by a few graph
    steps so the clocks stay where the timed region left them).  Per-op events carry a few us of overhead each and no PDL
    overlap, so their SUM overstates the graph step; the SHARES are what is used (cross-checked against the ncu launch list
    committed under profiles/)."""
    kinds = eng.b.prog.kinds
    shares, raw = [], []
    for _ in range(passes):
        eng.run(4)
        ms = eng.run_unet_timed()
        tot = sum(ms)
        by = {}
        for m, k in zip(ms, kinds):
            by[k] = by.get(k, 0.0) + m
        shares.append({k: v / tot for k, v in by.items()})
        raw.append(ms)
    med = {k: statistics.median(s.get(k, 0.0) for s in shares) for k in set(kinds)}
    med_ms = [statistics.median(r[i] for r in raw) for i in range(len(kinds))]
    return med, med_ms, statistics.median(sum(r) for r in raw)


def run_workload(wl, dev, rank, world, local_rank, steps, warmup, B=None, pair=None, detail=True, clock=False):
    """Runs `steps` timed jobs of workload `wl` (after `warmup`).  Returns a dict of measurements for this rank (timings are
    already the max over ranks)."""
    import torch
    import torch.distributed as dist

    from stablediffusion_b200 import _native as N
    from stablediffusion_b200.pipelines.sd_unified_pipeline import StableDiffusionUnifiedPipeline

    model_type, H, W, Bdef, sampler, sched_steps, guidance, strength = WORKLOADS[wl]
    B = B or Bdef
    n_exec = executed_steps(sched_steps, strength)
    model = build_model(wl, dev)
    pkw = {}
    if pair is not None:
        pkw = dict(pair_group=pair[0], pair_rank=pair[1], pair_exchange=pair[2])
    pipe = StableDiffusionUnifiedPipeline(do_cfg=True, device=str(dev), **pkw)
    # every rank owns an independent prompt shard; the ranks of a cond/uncond pair share theirs
    shard = rank // 2 if pair is not None else rank
    inp = synth_inputs(model_type, B, H, W, seed=1 + shard, image=strength is not None)
    host = {k: v.pin_memory() for k, v in inp.items()}
    devin = {k: v.to(dev) for k, v in inp.items()}
    job = make_job(wl, model, pipe, dev, B)

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
        loop_ms = pipe.last_engine.last_loop_ms()       # the LAST job's step loop, timed on the engine's stream
        t = torch.tensor([dev_s, wall, loop_ms], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        barrier()
        return t[0].item(), t[1].item(), t[2].item(), out

    for _ in range(max(warmup, 1)):
        job(devin, False)
    clk = ClockSampler(local_rank) if clock else contextlib.nullcontext()
    with clk:
        dev_s, _, loop_ms, _ = timed(devin, False, steps)
    job(host, True)
    _, e2e_wall, _, himg = timed(host, True, steps)

    eng = pipe.last_engine
    vae_prog = [v for k, v in model.vae._programs.items() if k[0] == "dec"][0]["b"].prog
    enc_launches = sum(v["b"].prog.num_launches for k, v in model.vae._programs.items() if k[0] == "enc")
    launches_per_job = 1 + n_exec * eng.launches_per_step + vae_prog.num_launches + enc_launches
    units = world // 2 if pair is not None else world   # independent image shards
    images = B * units * steps
    pk = peaks()
    tflop = job_tflop(wl, B)
    step_ms = loop_ms / n_exec
    unet_gf = UNET_GFLOP.get((model_type, H))
    res = dict(wl=wl, B=B, n_exec=n_exec, value=images / dev_s, e2e=images / e2e_wall, ms_per_job=1000.0 * dev_s / steps,
               unet_step_ms=step_ms, unet_steps_per_s=n_exec * units * steps / dev_s,
               unet_step_tensor_frac=(2 * B * unet_gf / 1000.0 / (step_ms * 1e-3) / pk["tf"]) if unet_gf else None,
               job_tflop=tflop, tensor_frac_whole_job=(tflop * steps / dev_s / pk["tf"]) if tflop else None,
               launches_per_step=eng.launches_per_step, launches_per_job=launches_per_job,
               h2d=sum(v.numel() * v.element_size() for v in host.values()), d2h=himg.numel() * himg.element_size(),
               clocks=clk.summary() if clock else None, eng=eng, model=model, pipe=pipe,
               exchange=getattr(eng, "exchange_kind", None))
    if detail and (rank == 0 or pair is not None):   # (the ranks of a cond/uncond pair step together: the exchange needs both)
        share, med_ms, sum_ms = step_breakdown(eng, N)
        kinds = eng.b.prog.kinds
        n_gemm = sum(1 for k in kinds if k == N.OP_GEMM)
        gemm_ms = step_ms * share.get(N.OP_GEMM, 0.0)
        att_ms = step_ms * share.get(N.OP_ATTENTION, 0.0)
        achieved = eng.gemm_flops / (gemm_ms * 1e-3) / 1e12
        traffic, traffic_src = None, None   # DRAM bytes per launch of the same kernel from the committed ncu capture
        tps = sorted(f for f in os.listdir(os.path.join(ROOT, "profiles")) if f.endswith("_gemm_traffic.json"))
        if wl == "sd15_512_b8_ddim50_cfg" and tps:   # the latest committed capture of the same command
            with open(os.path.join(ROOT, "profiles", tps[-1])) as f:
                td = json.load(f)
            traffic, traffic_src = td["dram_bytes_per_launch"], f"profiles/{tps[-1]} (ncu launch list of this workload, cold cache per launch)"
        res["roofline"] = {
            "bound": "tensor", "kernel": "gemm_tc_kernel (tcgen05 implicit GEMM, all conv/linear launches of one sampler step)",
            "achieved": achieved, "peak": pk["tf"], "unit": "TFLOP/s", "frac": achieved / pk["tf"], "traffic": traffic,
            "traffic_source": traffic_src, "algorithmic_flop_per_launch": eng.gemm_flops / n_gemm,
            "peak_source": f"{pk['src']} bf16_tflops_sustained", "launches": n_gemm, "avg_launch_ms": gemm_ms / n_gemm,
            "method": "graph step time inside the timed region x median share of the kernel over 5 per-op event passes",
            "step_ms_in_graph": step_ms, "step_ms_sum_of_per_op_events": sum_ms,
            "step_share_by_kind": {N.OP_NAMES[k]: round(v, 4) for k, v in sorted(share.items())},
            "step_ms_by_kind": {N.OP_NAMES[k]: round(step_ms * v, 4) for k, v in sorted(share.items())},
            "attention_tflops": (eng.attn_flops / (att_ms * 1e-3) / 1e12) if att_ms > 0 else None}
        # The attention kernels of this model are bound by the MUFU ex2 rate, not by the tensor pipe (head dim 40..160: one
        # exponential per score against 4*d tensor flops): 16 ex2 per clock and SM (measured, tools/micro/mufu.cu) at the SM clock
        # sampled during the timed region.  Reported for the self-attention launches of the largest sequence (the dominant kernel).
        att_idx = [i for i, k in enumerate(kinds) if k == N.OP_ATTENTION and i in eng.b.meta]
        if att_idx and res["clocks"] and res["clocks"].get("sm_mhz"):
            big = max(att_idx, key=lambda i: eng.b.meta[i]["Nq"] * eng.b.meta[i]["Nkv"])
            m = eng.b.meta[big]
            same = [i for i in att_idx if all(eng.b.meta[i][k] == m[k] for k in ("B", "heads", "Nq", "Nkv", "d"))]
            us = sum(med_ms[i] for i in same) / len(same) * 1e3 * (step_ms / sum_ms)
            exps = float(m["B"]) * m["heads"] * m["Nq"] * m["Nkv"]
            xu_peak = 16.0 * torch.cuda.get_device_properties(dev).multi_processor_count * res["clocks"]["sm_mhz"] * 1e6
            res["roofline"]["attention_xu"] = {
                "bound": "xu (MUFU ex2)", "kernel": f"attention_kernel, {m['B']}x{m['heads']} heads, {m['Nq']}x{m['Nkv']}, d={m['d']}",
                "launches_per_step": len(same), "us_per_launch_in_step": us, "exp_per_launch": exps, "achieved_gexp_s": exps / (us * 1e-6) / 1e9,
                "peak_gexp_s": xu_peak / 1e9, "frac": exps / (us * 1e-6) / xu_peak, "tflops": m["flops"] / (us * 1e-6) / 1e12,
                "peak_source": "16 ex2/clk/SM x SMs x median SM clock under load (clocks.sm_mhz)"}
        # HBM-bound kernels: algorithmic bytes = one read + one write of the tensor in bf16
        scale = step_ms / sum_ms
        hb = {}
        for kind, name in ((N.OP_GROUPNORM, "groupnorm_silu"), (N.OP_LAYERNORM, "layernorm")):
            idx = [i for i, k in enumerate(kinds) if k == kind]
            nbytes = sum(eng.b.meta[i]["bytes"] for i in idx if i in eng.b.meta)
            t_ms = step_ms * share.get(kind, 0.0)
            big = max(idx, key=lambda i: eng.b.meta.get(i, {}).get("bytes", 0)) if idx else None
            ent = {"launches_per_step": len(idx) * (1 if kind == N.OP_LAYERNORM else 2), "algorithmic_bytes_per_step": nbytes, "in_step_ms": t_ms,
                   "achieved_gbs": nbytes / (t_ms * 1e-3) / 1e9 if t_ms > 0 else None, "peak_gbs": pk["hbm"],
                   "frac": nbytes / (t_ms * 1e-3) / 1e9 / pk["hbm"] if t_ms > 0 else None}
            if big is not None:
                m = eng.b.meta[big]
                us = med_ms[big] * 1e3 * scale
                ent["largest_instance"] = {"shape": {k: v for k, v in m.items() if k not in ("op", "bytes")}, "bytes": m["bytes"], "us": us,
                                           "gbs": m["bytes"] / (us * 1e-6) / 1e9, "frac": m["bytes"] / (us * 1e-6) / 1e9 / pk["hbm"]}
            hb[name] = ent
        res["roofline_hbm"] = hb
    return res


def gpu_yardstick(dev, B_unet=16, h=64, iters=5):
    """torch-eager bf16 of the restated UNet modules (cuDNN 9 / cuBLASLt / SDPA flash on sm_100) at the headline UNet batch:
    ms per UNet forward on the same GPU.  A same-box library number to beat — NOT the reference, and not
    part of the product: the oracle modules are only timed here."""
    import torch

    from oracle.unet import SD15_UNET, UNet2DConditionOracle

    with torch.device(dev):
        net = UNet2DConditionOracle(**SD15_UNET).to(torch.bfloat16).eval()
    x = torch.randn(B_unet, 4, h, h, device=dev, dtype=torch.bfloat16)
    ctx = torch.randn(B_unet, 77, 768, device=dev, dtype=torch.bfloat16)
    t = torch.tensor(500.0, device=dev)
    with torch.no_grad():
        for _ in range(3):
            net(x, t, ctx)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            net(x, t, ctx)
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    del net
    torch.cuda.empty_cache()
    return {"what": "torch-eager bf16 forward of the restated SD1.5 UNet (NCHW, cuDNN/cuBLASLt/SDPA), UNet batch %d at 64x64" % B_unet,
            "ms_per_unet_forward": ms, "torch": torch.__version__, "role": "same-GPU library yardstick, not the reference"}


def run_b200(args, wl, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    model_type, H, W, Bdef, sampler, sched_steps, guidance, strength = WORKLOADS[wl]
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a B200: stablediffusion_b200 has no CPU fallback (use --impl reference for the CPU path)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    pair = None
    if args.pair_split:
        from stablediffusion_b200.distributed import make_pair_group

        if world % 2:
            raise SystemExit("--pair-split needs an even number of GPUs")
        grp, pr = make_pair_group()
        pair = (grp, pr, args.pair_exchange)
    B = args.batch or Bdef
    scaling = "weak"
    if args.strong:   # strong scaling: the GLOBAL batch is the workload's batch, split over the image shards
        units = world // 2 if pair is not None else world
        if B % units:
            raise SystemExit(f"--strong: global batch {B} does not divide over {units} shards")
        B, scaling = B // units, "strong"
    r = run_workload(wl, dev, rank, world, local_rank, args.steps, args.warmup, B=B, pair=pair, detail=True, clock=True)
    pk = peaks()
    extras, yard = None, None
    if rank == 0 and world == 1 and not args.no_extras and pair is None:
        # keep the headline engine's memory only as long as needed
        eng_launches = r["launches_per_step"]
        del r["eng"], r["model"], r["pipe"]
        torch.cuda.empty_cache()
        extras = {}
        for w2 in EXTRA_WORKLOADS:
            if w2 == wl:
                continue
            x = run_workload(w2, dev, 0, 1, local_rank, 3, 2, detail=False)
            m2, H2, W2, B2, s2, n2, g2, st2 = WORKLOADS[w2]
            extras[w2] = {"images_per_s": x["value"], "e2e_images_per_s": x["e2e"], "ms_per_job": x["ms_per_job"], "jobs_timed": 3,
                          "batch": B2, "sampler_steps": x["n_exec"], "unet_step_ms": x["unet_step_ms"],
                          "unet_step_tensor_frac": x["unet_step_tensor_frac"], "tensor_frac_whole_job": x["tensor_frac_whole_job"],
                          "launches_per_sampler_step": x["launches_per_step"], "algorithmic_tflop_per_job": x["job_tflop"]}
            del x
            torch.cuda.empty_cache()
        if not args.no_yardstick:
            try:
                yard = gpu_yardstick(dev)
                yard["ours_ms_per_unet_forward"] = r["unet_step_ms"] if wl == "sd15_512_b8_ddim50_cfg" and B == 8 else None
            except Exception as e:  # a yardstick must never take the bench line down
                yard = {"error": f"{type(e).__name__}: {e}"}
    if rank == 0:
        cb = None
        if world == 1 and not args.no_cpu_baseline:
            cb = cpu_baseline_for(wl)
        n_exec = r["n_exec"]
        par = f"dp{world}" if pair is None else f"dp{world // 2} x cond/uncond pair split ({r['exchange']} exchange)"
        line = {"metric": "images/s", "value": r["value"], "unit": "images/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": r["ms_per_job"], "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
                "dtype": "bf16", "data": "synthetic prompts (random embeddings), random-init weights of the named architecture",
                "config": {"workload": wl, "model": model_type, "resolution": f"{H}x{W}", "batch_per_gpu": r["B"], "scheduler": sampler,
                           "sampler_steps": n_exec, "guidance_scale": guidance, "cfg": True, "parallelism": par,
                           "l2": "working set (1.7 GB weights + activations) exceeds the 126 MB L2; no explicit flush"},
                "unet_steps_per_s": r["unet_steps_per_s"], "unet_step_ms": r["unet_step_ms"],
                "unet_step_tensor_frac": r["unet_step_tensor_frac"],
                "algorithmic_tflop_per_job": r["job_tflop"], "tensor_frac_whole_job": r["tensor_frac_whole_job"],
                "e2e": {"value": r["e2e"], "unit": "images/s", "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h"]},
                "gpu_launches": r["launches_per_job"] * args.steps, "launches_per_sampler_step": r["launches_per_step"],
                "clocks": r["clocks"], "roofline": r.get("roofline"), "roofline_hbm": r.get("roofline_hbm"),
                "extra_workloads": extras, "gpu_yardstick": yard, "cpu_baseline": cb}
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
    ap.add_argument("--batch", type=int, default=0, help="images per GPU (with --strong: the global batch); 0 = the workload's")
    ap.add_argument("--strong", action="store_true", help="strong scaling: fixed global batch split over the GPUs")
    ap.add_argument("--pair-split", action="store_true", help="cond/uncond split over GPU pairs (BASELINE config 5)")
    ap.add_argument("--pair-exchange", default="auto", choices=["auto", "peer", "nccl"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip extra_workloads (configs 1 / 3 / 4)")
    ap.add_argument("--no-yardstick", action="store_true")
    ap.add_argument("--no-config1", action="store_true", help="reference arm: skip the full timing of config 1")
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
