"""Request batching across jobs.

The reference's worker runs one `pipeline(model, **inference_config)` per job (runpod-worker/handler_logic.py:174-187), so
the UNet mostly sees batch 1-2, where a sampler step is launch-latency bound (5.0 ms per step for 1 image, 17.4 ms for 8 on
a B200: 3.5x the throughput).  `RequestBatcher` queues jobs and merges those that can share one call — same resolution,
sampler, step count, guidance scale, clip-skip, text-to-image mode — into one pipeline call of up to `max_batch` images.

A merged job gives every request the result it would have got alone: the reference seeds ONE generator per call and draws the
whole batch's noise from it (pipelines/sd_unified_pipeline.py:771-781), so each request's latents are drawn here from its own
seed, exactly as its solo call would, and handed to the pipeline through `latents=` (which only scales them by
init_noise_sigma, :783-785).  Image-to-image and inpainting jobs are not merged (their noise is drawn after the VAE
posterior, from the same generator: order D7) and run alone.

Pure host logic: the pipeline object is duck-typed (`pipeline(model, **kw) -> tensor [N, 3, H, W]`).
"""
from __future__ import annotations

from concurrent.futures import Future
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional, Tuple

import torch

_MERGE_KEYS = ("height", "width", "num_inference_steps", "guidance_scale", "clip_skip", "denoising_end", "guidance_rescale")
_SOLO_KEYS = ("image", "mask_image", "masked_image_latents", "latents", "cross_attention_kwargs", "denoising_start")


@dataclass
class _Request:
    config: Dict[str, Any]
    n_images: int
    future: Future = field(default_factory=Future)


def _as_list(p):
    return [p] if isinstance(p, str) else list(p)


class RequestBatcher:
    def __init__(self, pipeline, model, max_batch: int = 8):
        if max_batch < 1:
            raise ValueError("max_batch must be >= 1")
        self.pipeline, self.model, self.max_batch = pipeline, model, int(max_batch)
        self._queue: List[_Request] = []
        self.calls = 0  # pipeline calls issued so far

    # ---- queueing ---------------------------------------------------------------------------------------
    def submit(self, **inference_config) -> Future:
        """Queue one job (the keyword arguments of a pipeline call); the future resolves to its image tensor."""
        cfg = dict(inference_config)
        if "prompt_embeds" in cfg:
            n = cfg["prompt_embeds"].shape[0]
        elif cfg.get("prompt") is not None:
            n = len(_as_list(cfg["prompt"])) * int(cfg.get("num_images_per_prompt") or 1)
        else:
            raise ValueError("`prompt` (str or list) or `prompt_embeds` is required")
        req = _Request(cfg, n)
        self._queue.append(req)
        return req.future

    def _key(self, r: _Request) -> Optional[Tuple]:
        c = r.config
        if any(c.get(k) is not None for k in _SOLO_KEYS) or r.n_images > self.max_batch:
            return None
        if ("prompt_embeds" in c) != ("prompt" not in c or c.get("prompt") is None):
            return None
        kind = "embeds" if "prompt_embeds" in c else "text"
        if kind == "embeds" and ("pooled_prompt_embeds" in c):
            return None  # SDXL pre-encoded jobs: run alone (keeps this helper small)
        # pre-encoded jobs with and without negative embeddings do not mix (a merged call needs all or none)
        has_neg = kind == "embeds" and c.get("negative_prompt_embeds") is not None
        return (kind, has_neg, getattr(self.model, "scheduler_name", None)) + tuple(c.get(k) for k in _MERGE_KEYS)

    def plan(self) -> List[List[_Request]]:
        """Greedy first-fit in arrival order: a batch takes later compatible jobs while they fit."""
        batches: List[List[_Request]] = []
        open_by_key: Dict[Tuple, List[_Request]] = {}
        for r in self._queue:
            k = self._key(r)
            if k is None:
                batches.append([r])
                continue
            cur = open_by_key.get(k)
            if cur is not None and sum(x.n_images for x in cur) + r.n_images <= self.max_batch:
                cur.append(r)
            else:
                cur = [r]
                open_by_key[k] = cur
                batches.append(cur)
        return batches

    # ---- execution --------------------------------------------------------------------------------------
    def _latents_for(self, r: _Request, device, dtype) -> torch.Tensor:
        c = r.config
        m = self.model
        f = m.vae_scale_factor
        h = (c.get("height") or m.base.config.sample_size * f) // f
        w = (c.get("width") or m.base.config.sample_size * f) // f
        gen = None
        if c.get("seed") is not None:
            gen = torch.Generator(device=device).manual_seed(int(c["seed"]))
        return torch.randn((r.n_images, m.base.config.in_channels, h, w), generator=gen, device=device, dtype=dtype)

    def _run_batch(self, batch: List[_Request]):
        if len(batch) == 1:
            self.calls += 1
            return [self.pipeline(self.model, **batch[0].config)]
        first = batch[0].config
        kw = {k: first.get(k) for k in _MERGE_KEYS if first.get(k) is not None}
        device = self.pipeline.device
        if "prompt_embeds" in first:
            kw["prompt_embeds"] = torch.cat([r.config["prompt_embeds"] for r in batch])
            if first.get("negative_prompt_embeds") is not None:
                kw["negative_prompt_embeds"] = torch.cat([r.config["negative_prompt_embeds"] for r in batch])
            dtype = kw["prompt_embeds"].dtype
        else:
            # the second text encoder's prompts (SDXL; the reference worker sets both, handler_logic.py:315-327) follow the
            # pipeline's own defaults per request: prompt_2 = prompt, negative_prompt_2 = negative_prompt (:600, :620-621)
            prompts, negs, prompts2, negs2 = [], [], [], []
            for r in batch:
                c = r.config
                per = int(c.get("num_images_per_prompt") or 1)
                p = _as_list(c["prompt"])

                def per_prompt(v, default):
                    v = v or default
                    return [v] * len(p) if isinstance(v, str) else list(v)
                npmt = per_prompt(c.get("negative_prompt"), "")
                p2 = per_prompt(c.get("prompt_2"), p)
                n2 = per_prompt(c.get("negative_prompt_2"), npmt)
                if not (len(npmt) == len(p2) == len(n2) == len(p)):
                    raise ValueError("prompt, prompt_2, negative_prompt and negative_prompt_2 of a job must have the same length")
                for a, b, a2, b2 in zip(p, npmt, p2, n2):   # num_images_per_prompt is unrolled so that one merged call serves all
                    prompts += [a] * per
                    negs += [b] * per
                    prompts2 += [a2] * per
                    negs2 += [b2] * per
            kw["prompt"], kw["negative_prompt"] = prompts, negs
            if any(r.config.get("prompt_2") is not None or r.config.get("negative_prompt_2") is not None for r in batch):
                kw["prompt_2"], kw["negative_prompt_2"] = prompts2, negs2
            dtype = getattr(self.model.base, "dtype", torch.bfloat16)
        kw["latents"] = torch.cat([self._latents_for(r, device, dtype) for r in batch])
        self.calls += 1
        out = self.pipeline(self.model, **kw)
        res, o = [], 0
        for r in batch:
            res.append(out[o:o + r.n_images])
            o += r.n_images
        return res

    def run_pending(self) -> int:
        """Runs everything queued; returns the number of pipeline calls it took."""
        before = self.calls
        batches, self._queue = self.plan(), []
        for batch in batches:
            try:
                for r, img in zip(batch, self._run_batch(batch)):
                    r.future.set_result(img)
            except Exception as e:  # a failing batch fails its own jobs only
                for r in batch:
                    if not r.future.done():
                        r.future.set_exception(e)
        return self.calls - before
