"""Request batching across jobs: host logic on CPU with a recording pipeline; on the GPU, merged jobs
reproduce the solo results."""
from types import SimpleNamespace

import pytest
import torch

from stablediffusion_b200.serving import RequestBatcher


class _FakePipe:
    device = torch.device("cpu")

    def __init__(self):
        self.calls = []

    def __call__(self, model, **kw):
        self.calls.append(kw)
        if "latents" in kw and kw["latents"] is not None:
            n = kw["latents"].shape[0]
            tag = kw["latents"].flatten(1).sum(1)   # the result depends on each image's own noise only
        else:
            n = len(kw["prompt"]) if isinstance(kw["prompt"], list) else 1
            tag = torch.zeros(n)
        return tag.view(n, 1, 1, 1).expand(n, 3, 2, 2).clone()


def _model():
    return SimpleNamespace(vae_scale_factor=8, scheduler_name="euler",
                           base=SimpleNamespace(config=SimpleNamespace(sample_size=64, in_channels=4), dtype=torch.float32))


def test_compatible_jobs_share_calls_and_keep_their_noise():
    pipe, model = _FakePipe(), _model()
    b = RequestBatcher(pipe, model, max_batch=4)
    common = dict(num_inference_steps=20, guidance_scale=7.5, height=64, width=64)
    futs = [b.submit(prompt=f"p{i}", seed=100 + i, **common) for i in range(5)]
    futs.append(b.submit(prompt=["a", "b"], seed=7, **common))                     # 2 images: joins the second batch
    futs.append(b.submit(prompt="hi-res", seed=1, num_inference_steps=20, guidance_scale=7.5, height=128, width=128))
    assert b.run_pending() == 3 and len(pipe.calls) == 3
    assert [len(c["prompt"]) for c in pipe.calls[:2]] == [4, 3] and pipe.calls[1]["prompt"] == ["p4", "a", "b"]
    # each job's result equals what its own solo call would draw: randn(generator(seed)) of its own shape
    for i, f in enumerate(futs[:5]):
        want = torch.randn((1, 4, 8, 8), generator=torch.Generator().manual_seed(100 + i)).sum()
        assert torch.allclose(f.result()[0, 0, 0, 0], want)
    want = torch.randn((2, 4, 8, 8), generator=torch.Generator().manual_seed(7)).flatten(1).sum(1)
    assert futs[5].result().shape[0] == 2 and torch.allclose(futs[5].result()[:, 0, 0, 0], want)
    assert futs[6].result().shape[0] == 1 and "latents" not in pipe.calls[2]      # ran alone, untouched config


def test_second_encoder_prompts_are_forwarded_per_request():
    """SDXL jobs carry prompt_2 / negative_prompt_2 (the reference worker sets both, handler_logic.py:315-327): a merged call
    hands every request its own, with the pipeline's defaults (prompt_2 = prompt, negative_prompt_2 = negative_prompt)."""
    pipe = _FakePipe()
    rb = RequestBatcher(pipe, _model(), max_batch=8)
    rb.submit(prompt="a", prompt_2="a2", negative_prompt="na", seed=1)
    rb.submit(prompt=["b", "c"], negative_prompt_2=["nb2", "nc2"], seed=2)
    assert rb.run_pending() == 1
    kw = pipe.calls[0]
    assert kw["prompt"] == ["a", "b", "c"] and kw["prompt_2"] == ["a2", "b", "c"]
    assert kw["negative_prompt"] == ["na", "", ""] and kw["negative_prompt_2"] == ["na", "nb2", "nc2"]
    pipe.calls.clear()
    rb.submit(prompt="a", seed=1)
    rb.submit(prompt="b", seed=2)
    rb.run_pending()
    assert "prompt_2" not in pipe.calls[0]          # SD1.5-style jobs: nothing extra is passed
    # pre-encoded jobs with and without negative embeddings never share a call
    pipe.calls.clear()
    e = torch.zeros(1, 77, 8)
    rb.submit(prompt_embeds=e, negative_prompt_embeds=e, seed=1)
    rb.submit(prompt_embeds=e, seed=2)
    assert rb.run_pending() == 2


def test_incompatible_jobs_run_alone():
    pipe, model = _FakePipe(), _model()
    b = RequestBatcher(pipe, model, max_batch=8)
    b.submit(prompt="x", num_inference_steps=20, guidance_scale=7.5)
    b.submit(prompt="y", num_inference_steps=30, guidance_scale=7.5)               # other step count
    b.submit(prompt="z", num_inference_steps=20, guidance_scale=5.0)               # other guidance
    b.submit(prompt="i2i", num_inference_steps=20, guidance_scale=7.5, image=torch.zeros(1, 3, 8, 8), strength=0.5)
    b.submit(prompt="w", num_inference_steps=20, guidance_scale=7.5)               # merges with the first
    assert b.run_pending() == 4
    assert pipe.calls[0]["prompt"] == ["x", "w"] and pipe.calls[3]["prompt"] == "i2i" and "image" in pipe.calls[3]
    with pytest.raises(ValueError):
        b.submit(num_inference_steps=3)


def test_failure_is_confined_to_its_batch():
    class Boom(_FakePipe):
        def __call__(self, model, **kw):
            if kw.get("num_inference_steps") == 13:
                raise RuntimeError("boom")
            return super().__call__(model, **kw)
    b = RequestBatcher(Boom(), _model(), max_batch=2)
    ok = b.submit(prompt="a", num_inference_steps=20)
    bad = b.submit(prompt="b", num_inference_steps=13)
    b.run_pending()
    assert ok.result().shape[0] == 1
    with pytest.raises(RuntimeError):
        bad.result()


@pytest.mark.gpu
def test_merged_jobs_match_solo_jobs(cuda_device):
    from stablediffusion_b200.models.stable_diffusion import SDModelWrapper
    from stablediffusion_b200.pipelines.sd_unified_pipeline import StableDiffusionUnifiedPipeline
    from tests.test_models_gpu import SMALL_UNET, _make_unet, _make_vae, _rel_l2

    dev = cuda_device
    _, unet, _ = _make_unet(dev, **SMALL_UNET)
    _, vae, _ = _make_vae(dev)
    model = SDModelWrapper(vae=vae, base=unet, device=str(dev))
    pipe = StableDiffusionUnifiedPipeline(do_cfg=True, device=str(dev), output_type="latents")
    g = torch.Generator().manual_seed(3)
    jobs = []
    for i in range(3):
        jobs.append(dict(prompt_embeds=torch.randn(1, 77, 64, generator=g).to(dev, torch.bfloat16),
                         negative_prompt_embeds=torch.randn(1, 77, 64, generator=g).to(dev, torch.bfloat16),
                         num_inference_steps=6, guidance_scale=6.0, height=128, width=128, seed=40 + i))
    solo = [pipe(model, **j).float().cpu() for j in jobs]
    b = RequestBatcher(pipe, model, max_batch=8)
    futs = [b.submit(**j) for j in jobs]
    assert b.run_pending() == 1
    for f, s in zip(futs, solo):
        # same noise, same prompts; the UNet batch differs, so GEMM tilings (bf16 rounding points) may: trajectory tolerance
        assert _rel_l2(f.result().float().cpu(), s) <= 5e-2
