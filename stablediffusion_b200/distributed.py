"""Multi-GPU execution of the hot path: one process per GPU (torchrun), independent image / prompt shards.

The reference never shards inference (one global model per worker process, runpod-worker/rp_handler.py:15,36);
each image's trajectory depends only on its own latent, prompt embedding and seed, so the path partitions with
NO data-path collective:
  * data parallel: rank r takes items [r*B/G, (r+1)*B/G) of the batch; to stay bit-identical to a single-GPU run
    every rank draws the FULL (B,4,h,w) noise tensor from the one seeded generator (reference :771-781 seeds one
    generator for the whole batch) and slices its shard; weights are replicated; one all-gather of the output
    images at the end (optional).
  * optional cond/uncond split on GPU pairs (BASELINE config 5): ranks (2p, 2p+1) hold identical latents, run the
    uncond / cond half of the CFG batch, and exchange the eps halves with ONE NCCL all-gather per sampler step
    (message B*4*h*w*2 bytes) — see engine.DenoiseEngine.
Host-side logic here is backend-agnostic (tested with gloo on CPU); tensors that cross ranks are torch tensors.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def shard_range(n_items: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced shard [lo, hi) of n_items for `rank` (first n_items % world ranks get one extra)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_list(items: Sequence, rank: int, world: int) -> list:
    lo, hi = shard_range(len(items), rank, world)
    return list(items[lo:hi])


def sharded_noise(shape: Sequence[int], seed: int, rank: int, world: int, device, dtype=torch.float32) -> torch.Tensor:
    """This rank's rows of the full-batch noise a single-GPU run would draw (reference :771-781)."""
    gen = torch.Generator(device=device).manual_seed(int(seed))
    full = torch.randn(tuple(shape), generator=gen, device=device, dtype=dtype)
    lo, hi = shard_range(shape[0], rank, world)
    return full[lo:hi].contiguous()


def gather_batches(local: torch.Tensor, sizes: Optional[List[int]] = None, group=None) -> torch.Tensor:
    """Concatenate per-rank batches (possibly ragged along dim 0) on every rank."""
    world = dist.get_world_size(group)
    if world == 1:
        return local
    if sizes is None:
        n = torch.tensor([local.shape[0]], device=local.device, dtype=torch.int64)
        all_n = [torch.zeros_like(n) for _ in range(world)]
        dist.all_gather(all_n, n, group=group)
        sizes = [int(x.item()) for x in all_n]
    mx = max(sizes)
    pad = torch.zeros((mx,) + tuple(local.shape[1:]), device=local.device, dtype=local.dtype)
    pad[: local.shape[0]] = local
    outs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(outs, pad, group=group)
    return torch.cat([o[:s] for o, s in zip(outs, sizes)], dim=0)


def make_pair_group(rank: Optional[int] = None, world: Optional[int] = None):
    """Process group of this rank's cond/uncond pair (2p, 2p+1) and the rank inside the pair.
    Every rank must call this (new_group is collective over the default group)."""
    rank = dist.get_rank() if rank is None else rank
    world = dist.get_world_size() if world is None else world
    if world % 2 != 0:
        raise ValueError("the cond/uncond split needs an even number of ranks")
    mine = None
    for p in range(world // 2):
        g = dist.new_group([2 * p, 2 * p + 1])
        if rank // 2 == p:
            mine = g
    return mine, rank % 2


def generate_data_parallel(pipe, model, prompt_embeds, negative_prompt_embeds, seed: int, gather: bool = True, **kw):
    """Run `pipe` on this rank's shard of a global batch of pre-encoded prompts; returns the gathered images
    (every rank) or the local shard.  Results equal a single-GPU run of the whole batch item for item."""
    rank, world = (dist.get_rank(), dist.get_world_size()) if dist.is_initialized() else (0, 1)
    B = prompt_embeds.shape[0]
    lo, hi = shard_range(B, rank, world)
    h = (kw.get("height") or model.base.config.sample_size * model.vae_scale_factor) // model.vae_scale_factor
    w = (kw.get("width") or model.base.config.sample_size * model.vae_scale_factor) // model.vae_scale_factor
    lat = sharded_noise((B, model.base.config.in_channels, h, w), seed, rank, world, pipe.device, prompt_embeds.dtype)
    if hi == lo:
        local = torch.zeros((0, 3, h * 8, w * 8), device=pipe.device, dtype=prompt_embeds.dtype)
    else:
        local = pipe(model, prompt_embeds=prompt_embeds[lo:hi], negative_prompt_embeds=negative_prompt_embeds[lo:hi],
                     latents=lat, **kw)
    if gather and world > 1:
        return gather_batches(local)
    return local
