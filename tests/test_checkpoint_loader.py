"""`SDModelWrapper.load_hf_checkpoint` (reference models/stable_diffusion.py:106-156): an HF-layout folder — unet/, vae/,
text_encoder/, tokenizer/, scheduler/ with fp16-variant safetensors, i.e. what scripts/convert_from_A1111.py:1381-1386
writes — is read, converted and (on the GPU) driven from a text prompt to an image through the reference's call sequence."""
import json
import os

import pytest
import torch

from stablediffusion_b200.config import scheduler_config, unet_config, vae_config
from stablediffusion_b200.random_init import random_state_dict
from stablediffusion_b200.unet import unet_param_shapes
from stablediffusion_b200.vae import vae_param_shapes

SMALL_UNET = dict(block_out_channels=(64, 128, 128), layers_per_block=1, cross_attention_dim=64, attention_head_dim=2,
                  down_block_types=("CrossAttnDownBlock2D", "CrossAttnDownBlock2D", "DownBlock2D"),
                  up_block_types=("UpBlock2D", "CrossAttnUpBlock2D", "CrossAttnUpBlock2D"), sample_size=16)
SMALL_VAE = dict(block_out_channels=(32, 64, 64, 64), layers_per_block=1)
CLIP_CFG = dict(vocab_size=515, hidden_size=64, intermediate_size=128, num_hidden_layers=2, num_attention_heads=2,
                max_position_embeddings=77, hidden_act="quick_gelu", bos_token_id=512, eos_token_id=513, pad_token_id=513)


def _write_tokenizer(folder):
    """A byte-level CLIP BPE vocabulary without merges: every byte symbol, bare and word-final, plus the two specials."""
    os.makedirs(folder, exist_ok=True)
    bs = list(range(ord("!"), ord("~") + 1)) + list(range(ord("\xa1"), ord("\xac") + 1)) + list(range(ord("\xae"), ord("\xff") + 1))
    cs, n = bs[:], 0
    for b in range(256):          # GPT-2 / CLIP byte-to-unicode table: printable bytes map to themselves, the rest to 256+
        if b not in bs:
            bs.append(b)
            cs.append(256 + n)
            n += 1
    syms = [chr(c) for c in cs]
    vocab = {s: i for i, s in enumerate(syms)}
    vocab.update({s + "</w>": 256 + i for i, s in enumerate(syms)})
    vocab["<|startoftext|>"], vocab["<|endoftext|>"] = 512, 513
    vocab["zz</w>"] = 514
    json.dump(vocab, open(os.path.join(folder, "vocab.json"), "w"))
    open(os.path.join(folder, "merges.txt"), "w").write("#version: 0.2\nz z</w>\n")
    json.dump({"model_max_length": 77, "bos_token": "<|startoftext|>", "eos_token": "<|endoftext|>", "unk_token": "<|endoftext|>",
               "pad_token": "<|endoftext|>", "tokenizer_class": "CLIPTokenizer"}, open(os.path.join(folder, "tokenizer_config.json"), "w"))


def make_checkpoint(root):
    from safetensors.torch import save_file
    from transformers import CLIPTextConfig, CLIPTextModel

    ucfg, vcfg = unet_config(**SMALL_UNET), vae_config(**SMALL_VAE)
    usd = random_state_dict(unet_param_shapes(ucfg), seed=0, dtype=torch.float16)
    vsd = random_state_dict(vae_param_shapes(vcfg, True, True), seed=1, dtype=torch.float16)
    for sub, cfg, sd, cls in (("unet", ucfg, usd, "UNet2DConditionModel"), ("vae", vcfg, vsd, "AutoencoderKL")):
        os.makedirs(os.path.join(root, sub))
        c = {k: (list(v) if isinstance(v, tuple) else v) for k, v in cfg.items()}
        c.update(_class_name=cls, _diffusers_version="0.27.2")
        json.dump(c, open(os.path.join(root, sub, "config.json"), "w"))
        save_file(sd, os.path.join(root, sub, "diffusion_pytorch_model.fp16.safetensors"))   # the fp16 variant name (:113-121)
    os.makedirs(os.path.join(root, "scheduler"))
    sc = dict(scheduler_config("sd15"))
    sc.pop("trained_betas")
    sc.update(_class_name="PNDMScheduler", _diffusers_version="0.27.2")   # SD1.5's stock scheduler_config.json
    json.dump(sc, open(os.path.join(root, "scheduler", "scheduler_config.json"), "w"))
    torch.manual_seed(0)
    CLIPTextModel(CLIPTextConfig(**CLIP_CFG)).save_pretrained(os.path.join(root, "text_encoder"))
    _write_tokenizer(os.path.join(root, "tokenizer"))
    return ucfg, vcfg, usd, vsd


def test_load_hf_checkpoint_reads_the_folder_layout(tmp_path):
    from stablediffusion_b200.models.stable_diffusion import SDModelWrapper
    from stablediffusion_b200.schedulers import B200DDIMScheduler, B200EulerDiscreteScheduler

    root = str(tmp_path / "ckpt")
    os.makedirs(root)
    ucfg, vcfg, usd, vsd = make_checkpoint(root)
    m = SDModelWrapper(ckpt_path=root, device="cpu")
    assert m.path == root and m.type == "sd15" and m.scheduler_name == "euler"
    assert tuple(m.base.config.block_out_channels) == (64, 128, 128) and m.base.config.cross_attention_dim == 64
    assert set(m.base._master) == set(usd) and all(torch.equal(m.base._master[k], usd[k]) for k in usd)
    assert set(m.vae._master) == set(vsd) and m.vae.config.scaling_factor == 0.18215 and m.vae_scale_factor == 8
    # default scheduler: Euler built from the checkpoint's (PNDM-style) config (:135-138): class defaults fill the rest
    assert isinstance(m.scheduler, B200EulerDiscreteScheduler)
    assert m.scheduler.config.timestep_spacing == "linspace" and m.scheduler.config.steps_offset == 1
    assert m.scheduler.config.beta_schedule == "scaled_linear" and m.scheduler.config.skip_prk_steps is True
    m.set_scheduler("DDIM")
    assert isinstance(m.scheduler, B200DDIMScheduler) and m.scheduler.config.timestep_spacing == "leading"
    m.scheduler.set_timesteps(50)
    assert m.scheduler.timesteps.tolist()[:2] == [981, 961]
    ids = m.tokenizer(["zz a"], padding="max_length", max_length=77, truncation=True, return_tensors="pt").input_ids
    assert ids.shape == (1, 77) and ids[0, 0].item() == 512 and ids[0, 1].item() == 514
    m.load_hf_checkpoint(root)            # same path: early return (:107-108)
    with pytest.raises(FileNotFoundError):
        SDModelWrapper(ckpt_path=str(tmp_path / "missing"), device="cpu")
    with pytest.raises(FileNotFoundError, match="not a local folder"):
        SDModelWrapper(model_name="anything", model_type="sd15", device="cpu")   # hub ids need a network


@pytest.mark.gpu
def test_checkpoint_to_image_from_a_text_prompt(cuda_device, tmp_path):
    """load -> .to(cuda) (weights packed, CLIP becomes the native encoder) -> pipeline(model, prompt=...) -> image; the
    UNet the folder produced is the oracle's UNet on the same fp16 weights."""
    from oracle.unet import UNet2DConditionOracle
    from stablediffusion_b200.models.stable_diffusion import SDModelWrapper
    from stablediffusion_b200.pipelines.sd_unified_pipeline import StableDiffusionUnifiedPipeline

    dev = cuda_device
    root = str(tmp_path / "ckpt")
    os.makedirs(root)
    ucfg, vcfg, usd, vsd = make_checkpoint(root)
    m = SDModelWrapper(ckpt_path=root, device=str(dev))
    pipe = StableDiffusionUnifiedPipeline(do_cfg=True, device=str(dev))
    img = pipe(m, prompt=["zz a b", "c"], negative_prompt="zz", num_inference_steps=6, guidance_scale=5.0, seed=3, height=128, width=128)
    torch.cuda.synchronize()
    assert img.shape == (2, 3, 128, 128) and torch.isfinite(img.float()).all()
    img2 = pipe(m, prompt=["zz a b", "c"], negative_prompt="zz", num_inference_steps=6, guidance_scale=5.0, seed=3, height=128, width=128)
    assert torch.equal(img, img2)          # same seed, same result
    oracle = UNet2DConditionOracle(**{k: SMALL_UNET[k] for k in SMALL_UNET if k != "sample_size"}).eval()
    oracle.load_state_dict({k: v.to(torch.bfloat16).float() for k, v in usd.items()}, strict=True)
    g = torch.Generator().manual_seed(1)
    x = torch.randn(2, 4, 16, 16, generator=g).to(torch.bfloat16).float()
    ctx = torch.randn(2, 77, 64, generator=g).to(torch.bfloat16)
    out = m.base(x.to(dev), torch.tensor(300.0), ctx.to(dev), return_dict=False)[0].float().cpu()
    with torch.no_grad():
        ref = oracle(x, torch.tensor(300.0), ctx.float())
    rel = ((out - ref).norm() / ref.norm()).item()
    assert rel <= 2e-2, rel
