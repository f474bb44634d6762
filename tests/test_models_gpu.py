"""Model-level parity on the B200: B200UNet / B200VAE / fused loop (all through the C ABI) vs the fp32 oracle
on identical random-init weights (bf16-rounded), identical latents and context.

Tolerances: per-step UNet eps rel-L2 <= 2e-2 and max-abs <= 0.1*max|eps|; the same quantities of
torch-eager bf16 running the oracle modules on the same GPU are printed beside ours as the yardstick.
"""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

SMALL_UNET = dict(block_out_channels=(64, 128, 128), layers_per_block=1, cross_attention_dim=64, attention_head_dim=2,
                  down_block_types=("CrossAttnDownBlock2D", "CrossAttnDownBlock2D", "DownBlock2D"),
                  up_block_types=("UpBlock2D", "CrossAttnUpBlock2D", "CrossAttnUpBlock2D"), sample_size=16)


def _rel_l2(a, b):
    a, b = a.float(), b.float()
    return ((a - b).pow(2).sum().sqrt() / (b.pow(2).sum().sqrt() + 1e-12)).item()


def _bf16_round_sd(sd):
    return {k: v.to(torch.bfloat16).to(torch.float32) for k, v in sd.items()}


def _make_unet(dev, model_type="sd15", seed=0, **over):
    from oracle.unet import UNet2DConditionOracle
    from stablediffusion_b200.config import unet_config
    from stablediffusion_b200.random_init import random_state_dict
    from stablediffusion_b200.unet import B200UNet, unet_param_shapes

    cfg = unet_config(model_type, **over)
    sd = _bf16_round_sd(random_state_dict(unet_param_shapes(cfg), seed=seed))
    ours = B200UNet(cfg, sd, device=dev)
    okeys = {k: cfg[k] for k in ("in_channels", "out_channels", "block_out_channels", "layers_per_block", "down_block_types",
                                 "up_block_types", "attention_head_dim", "cross_attention_dim", "norm_num_groups", "norm_eps",
                                 "use_linear_projection", "transformer_layers_per_block", "flip_sin_to_cos", "freq_shift",
                                 "addition_embed_type", "addition_time_embed_dim", "projection_class_embeddings_input_dim")}
    oracle = UNet2DConditionOracle(**okeys)
    oracle.load_state_dict(sd, strict=True)
    return cfg, ours, oracle.eval()


def _unet_case(dev, cfg, ours, oracle, B, h, w, t, oracle_on_gpu):
    g = torch.Generator().manual_seed(11)
    x = torch.randn(B, 4, h, w, generator=g)
    ctx = torch.randn(B, 77, cfg["cross_attention_dim"], generator=g).to(torch.bfloat16).float()
    added = None
    if cfg.get("addition_embed_type") == "text_time":
        added = dict(text_embeds=torch.randn(B, 1280, generator=g).to(torch.bfloat16).float(),
                     time_ids=torch.tensor([[8.0 * h, 8.0 * w, 0, 0, 8.0 * h, 8.0 * w]]).repeat(B, 1))
    xb = x.to(torch.bfloat16).float()  # the UNet input is bf16 on our path; feed the oracle the same values
    odev = dev if oracle_on_gpu else torch.device("cpu")
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    oracle = oracle.to(odev)
    with torch.no_grad():
        oadd = None if added is None else {k: v.to(odev) for k, v in added.items()}
        ref = oracle(xb.to(odev), torch.tensor(t), ctx.to(odev), added_cond_kwargs=oadd).float().cpu()
        eager = None
        if oracle_on_gpu:
            ob = oracle.to(torch.bfloat16)
            badd = None if added is None else {k: v.to(odev, torch.bfloat16) for k, v in added.items()}
            eager = ob(xb.to(odev, torch.bfloat16), torch.tensor(t), ctx.to(odev, torch.bfloat16), added_cond_kwargs=badd).float().cpu()
            oracle.to(torch.float32)
    dadd = None if added is None else {k: v.to(dev) for k, v in added.items()}
    out = ours(xb.to(dev), torch.tensor(t), ctx.to(dev, torch.bfloat16), added_cond_kwargs=dadd, return_dict=False)[0]
    torch.cuda.synchronize()
    out = out.float().cpu()
    assert torch.isfinite(out).all()
    rl2 = _rel_l2(out, ref)
    mabs = (out - ref).abs().max().item() / ref.abs().max().item()
    msg = f"rel-L2 {rl2:.4g}, max-abs/max {mabs:.4g}"
    if eager is not None:
        msg += f" | torch-eager-bf16 rel-L2 {_rel_l2(eager, ref):.4g}"
    print("UNET PARITY", tuple(x.shape), "t=", t, msg)
    assert rl2 <= 2e-2, msg
    assert mabs <= 0.1, msg
    return rl2


def test_unet_small_vs_oracle_cpu(cuda_device):
    cfg, ours, oracle = _make_unet(cuda_device, **SMALL_UNET)
    _unet_case(cuda_device, cfg, ours, oracle, 2, 16, 16, 946.4211, oracle_on_gpu=False)
    _unet_case(cuda_device, cfg, ours, oracle, 3, 24, 16, 1.0, oracle_on_gpu=False)


def test_unet_sd15_512_vs_oracle(cuda_device):
    cfg, ours, oracle = _make_unet(cuda_device)
    _unet_case(cuda_device, cfg, ours, oracle, 2, 64, 64, 946.4211, oracle_on_gpu=True)
    print("launches per forward:", ours.num_launches)


def test_unet_sd15_768_vs_oracle(cuda_device):
    cfg, ours, oracle = _make_unet(cuda_device)
    _unet_case(cuda_device, cfg, ours, oracle, 2, 96, 96, 501.0, oracle_on_gpu=True)


def test_unet_sdxl_vs_oracle(cuda_device):
    cfg, ours, oracle = _make_unet(cuda_device, "sdxl")
    _unet_case(cuda_device, cfg, ours, oracle, 2, 64, 64, 958.0, oracle_on_gpu=True)
