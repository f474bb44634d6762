"""CLIP text encoders.

CPU: the oracle restatement is PINNED against fixtures produced by the installed `transformers` classes
(tests/golden/make_clip_golden.py).  GPU: the native encoder (stablediffusion_b200/clip.py, kernels through the C ABI)
against the oracle — bf16 kernels vs fp32 oracle, tolerance rel-L2 <= 2e-2 per hidden state (stated in SURVEY §8d).
"""
import os

import pytest
import torch

from oracle.clip import CLIPTextOracle

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _load(name):
    return torch.load(os.path.join(GOLD, name + ".pt"), weights_only=False)


def _rel(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a - b).norm() / b.norm().clamp_min(1e-12)).item()


@pytest.mark.parametrize("name", ["clip_l_tiny", "clip_g_tiny"])
def test_oracle_pinned_to_transformers(name):
    g = _load(name)
    o = CLIPTextOracle(g["config"], g["state_dict"], g["with_projection"])(g["input_ids"])
    assert len(o["hidden_states"]) == g["hidden_states"].shape[0]
    for i, h in enumerate(o["hidden_states"]):
        assert _rel(h, g["hidden_states"][i]) < 2e-6, i
    assert _rel(o["last_hidden_state"], g["last_hidden_state"]) < 2e-6
    if g["with_projection"]:
        assert _rel(o["text_embeds"], g["text_embeds"]) < 2e-6
    else:
        assert _rel(o["pooler_output"], g["pooler_output"]) < 2e-6


def test_oracle_is_causal():
    g = _load("clip_l_tiny")
    orc = CLIPTextOracle(g["config"], g["state_dict"])
    ids = g["input_ids"].clone()
    a = orc(ids)["last_hidden_state"]
    ids[:, 50:] = 7  # changing the future must not change the past
    b = orc(ids)["last_hidden_state"]
    assert torch.equal(a[:, :50], b[:, :50]) and not torch.equal(a[:, 50:], b[:, 50:])


def _native(g, device):
    from stablediffusion_b200.clip import B200CLIPTextModel
    return B200CLIPTextModel.from_state_dict(g["config"], g["state_dict"], with_projection=g["with_projection"], device=device)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["clip_l_tiny", "clip_g_tiny"])
def test_native_matches_golden(cuda_device, name):
    g = _load(name)
    enc = _native(g, cuda_device)
    out = enc(g["input_ids"].to(cuda_device), output_hidden_states=True)
    assert len(out.hidden_states) == g["hidden_states"].shape[0]
    for i, h in enumerate(out.hidden_states):
        assert h.dtype == torch.bfloat16 and _rel(h, g["hidden_states"][i]) < 2e-2, (i, _rel(h, g["hidden_states"][i]))
    if g["with_projection"]:
        assert _rel(out[0], g["text_embeds"]) < 2e-2 and _rel(out.text_embeds, g["text_embeds"]) < 2e-2
        assert _rel(out.last_hidden_state, g["last_hidden_state"]) < 2e-2
    else:
        assert _rel(out[0], g["last_hidden_state"]) < 2e-2
        assert _rel(out.pooler_output, g["pooler_output"]) < 2e-2
    # the clip-skip path of the reference (pipelines/sd_unified_pipeline.py:605-609)
    skip = enc.text_model.final_layer_norm(out[-1][-2])
    want = CLIPTextOracle(g["config"], g["state_dict"]).final_layer_norm(g["hidden_states"][-2])
    assert _rel(skip, want) < 2e-2


@pytest.mark.gpu
@pytest.mark.parametrize("variant", ["clip_l", "clip_g"])
def test_native_full_size_matches_oracle(cuda_device, variant):
    """Real tower shapes (CLIP-L 12x768 / OpenCLIP-bigG 32x1280 cut to 4 layers to keep the CPU oracle quick)."""
    from stablediffusion_b200.clip import B200CLIPTextModel, random_clip_state_dict
    cfg = dict(vocab_size=49408, max_position_embeddings=77, layer_norm_eps=1e-5)
    if variant == "clip_l":
        cfg.update(hidden_size=768, intermediate_size=3072, num_hidden_layers=4, num_attention_heads=12,
                   hidden_act="quick_gelu", eos_token_id=2, projection_dim=768)
    else:
        cfg.update(hidden_size=1280, intermediate_size=5120, num_hidden_layers=4, num_attention_heads=20,
                   hidden_act="gelu", eos_token_id=2, projection_dim=1280)
    sd = random_clip_state_dict(cfg, with_projection=variant == "clip_g", seed=5)
    g = torch.Generator().manual_seed(9)
    ids = torch.randint(1000, 40000, (4, 77), generator=g)
    ids[:, 0] = 49406
    for b, n in enumerate((3, 20, 50, 76)):
        ids[b, n:] = 49407
    enc = B200CLIPTextModel.from_state_dict(cfg, sd, with_projection=variant == "clip_g", device=cuda_device)
    out = enc(ids.to(cuda_device), output_hidden_states=True)
    # the oracle sees the same bf16-rounded weights
    ref = CLIPTextOracle(cfg, {k: v.bfloat16().float() for k, v in sd.items()}, variant == "clip_g")(ids)
    for i, h in enumerate(out.hidden_states):
        assert _rel(h, ref["hidden_states"][i]) < 2e-2, (i, _rel(h, ref["hidden_states"][i]))
    assert _rel(out.last_hidden_state, ref["last_hidden_state"]) < 2e-2
    if variant == "clip_g":
        assert _rel(out.text_embeds, ref["text_embeds"]) < 2e-2
    else:
        assert _rel(out.pooler_output, ref["pooler_output"]) < 2e-2
    # same ids, second call: the cached program replays bit-identically
    again = enc(ids.to(cuda_device), output_hidden_states=True)
    assert torch.equal(again.last_hidden_state, out.last_hidden_state)


@pytest.mark.gpu
def test_wrapper_swaps_in_native_encoder_and_merges_lora(cuda_device):
    """SDModelWrapper.to(cuda) turns a transformers CLIPTextModel into the native encoder; `encode_prompt` (reference
    pipeline :532-719) then runs on it; a text-encoder weight change on the torch master copy reaches the device."""
    from types import SimpleNamespace

    from transformers import CLIPTextConfig, CLIPTextModel

    from stablediffusion_b200.clip import B200CLIPTextModel
    from stablediffusion_b200.models.stable_diffusion import SDModelWrapper
    from stablediffusion_b200.pipelines.sd_unified_pipeline import StableDiffusionUnifiedPipeline
    from tests.test_models_gpu import SMALL_UNET, _make_unet, _make_vae

    g = _load("clip_l_tiny")
    mod = CLIPTextModel(CLIPTextConfig(bos_token_id=0, pad_token_id=1, **g["config"])).eval()
    mod.load_state_dict(g["state_dict"], strict=False)

    class Tok:
        model_max_length = 77

        def __call__(self, texts, **kw):
            assert kw["padding"] == "max_length" and kw["max_length"] == 77 and kw["truncation"]
            return SimpleNamespace(input_ids=g["input_ids"][:len(texts)].clone())

    dev = cuda_device
    _, unet, _ = _make_unet(dev, **SMALL_UNET)
    _, vae, _ = _make_vae(dev)
    model = SDModelWrapper(vae=vae, base=unet, text_encoder=mod, tokenizer=Tok(), device=str(dev))
    model.to(str(dev))
    assert isinstance(model.text_encoder, B200CLIPTextModel)
    pipe = StableDiffusionUnifiedPipeline(do_cfg=True, device=str(dev), output_type="latents")
    pipe.model = model
    pe, ne, pooled, npooled = pipe.encode_prompt(["a", "b", "c"], negative_prompt="")
    assert pe.shape == (3, 77, 64) and ne.shape == (3, 77, 64) and pe.dtype == torch.bfloat16
    assert _rel(pe, g["last_hidden_state"]) < 2e-2
    skip = pipe.encode_prompt(["a", "b", "c"], clip_skip=1)[0]
    want = CLIPTextOracle(g["config"], g["state_dict"]).final_layer_norm(g["hidden_states"][-2])
    assert _rel(skip, want) < 2e-2
    # weight change on the master copy (what a text-encoder LoRA merge does) -> refresh -> different embeddings
    with torch.no_grad():
        mod.text_model.encoder.layers[0].mlp.fc1.weight.mul_(1.5)
    model.text_encoder.refresh_weights()
    sd2 = {k: v for k, v in mod.state_dict().items() if "position_ids" not in k}
    want2 = CLIPTextOracle(g["config"], sd2)(g["input_ids"])["last_hidden_state"]
    got2 = pipe.encode_prompt(["a", "b", "c"])[0]
    assert _rel(got2, want2) < 2e-2 and _rel(got2, g["last_hidden_state"]) > 5e-2
