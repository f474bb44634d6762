"""CPU tests of the host side: scheduler tables (bit-exact vs oracle and golden fixtures), the fused step-table
coefficients (emulated in fp32 exactly as the CUDA kernel evaluates them), LoRA parsing/merging, weight packing,
config semantics, pipeline index arithmetic."""
import math
import os

import pytest
import torch

from oracle import schedulers as OS
from stablediffusion_b200 import _native as N
from stablediffusion_b200.config import scheduler_config, unet_config
from stablediffusion_b200.schedulers import (B200DDIMScheduler, B200DPMSolverMultistepScheduler, B200EulerDiscreteScheduler,
                                             make_scheduler)

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CFG = dict(OS.SD15_SCHEDULER_CONFIG)


@pytest.mark.parametrize("name,n", [("euler", 20), ("euler", 30), ("euler", 50), ("DDIM", 20), ("DDIM", 50), ("DPM++ 2M", 20),
                                    ("DPM++ 2M", 50), ("DPM++ 2M Karras", 25)])
def test_timestep_tables_bit_exact(name, n):
    gold = torch.load(os.path.join(GOLDEN, "scheduler_tables.pt"))[f"{name}/{n}"]
    s = make_scheduler(name, B200EulerDiscreteScheduler(**CFG).config)
    s.set_timesteps(n)
    o = OS.make_scheduler(name, CFG)
    o.set_timesteps(n)
    for ref in (gold["timesteps"], o.timesteps):
        assert s.timesteps.dtype == ref.dtype and torch.equal(s.timesteps, ref)          # int64 exact / fp32 bit pattern
    if gold["sigmas"] is not None and hasattr(s, "sigmas"):
        assert torch.equal(s.sigmas, gold["sigmas"]) and torch.equal(s.sigmas, o.sigmas)  # fp32 bit-exact
    assert float(s.init_noise_sigma) == gold["init_noise_sigma"]


def test_sdxl_euler_leading_table():
    gold = torch.load(os.path.join(GOLDEN, "scheduler_tables.pt"))["sdxl-euler/30"]
    s = B200EulerDiscreteScheduler(**scheduler_config("sdxl"))
    s.set_timesteps(30)
    assert torch.equal(s.timesteps, gold["timesteps"]) and torch.equal(s.sigmas, gold["sigmas"])
    assert float(s.init_noise_sigma) == gold["init_noise_sigma"]


def _emulate_kernel(tab_row, x, eps, hist):
    """fp32 evaluation in the order of cfg_step_kernel (csrc/bandwidth.cu)."""
    K1, K2, CX, CE, C0, C1 = [tab_row[i] for i in (N.ST_K1, N.ST_K2, N.ST_CX, N.ST_CE, N.ST_C0, N.ST_C1)]
    x0 = (x - K1 * eps) * K2
    xn = CX * x + CE * eps + C0 * x0
    if hist is not None and C1 != 0:
        xn = xn + C1 * (x0 - hist)
    return xn, x0


@pytest.mark.parametrize("name,n", [("euler", 20), ("DDIM", 50), ("DPM++ 2M", 20), ("DPM++ 2M Karras", 10), ("DPM++ 2M", 1)])
def test_step_table_reproduces_oracle_trajectory(name, n):
    s = make_scheduler(name, B200EulerDiscreteScheduler(**CFG).config)
    s.set_timesteps(n)
    o = OS.make_scheduler(name, CFG)
    o.set_timesteps(n)
    g = torch.Generator().manual_seed(0)
    x = torch.randn(2, 4, 8, 8, generator=g) * float(o.init_noise_sigma)
    xo, hist = x.clone(), torch.zeros_like(x)
    for i, t in enumerate(o.timesteps):
        eps = torch.randn(2, 4, 8, 8, generator=g)
        assert s._table[i, N.ST_TIMESTEP].item() == float(t)
        scaled = o.scale_model_input(xo, t)
        torch.testing.assert_close(x * s._table[i, N.ST_IN_SCALE], scaled, rtol=2e-5, atol=2e-5)
        xo = o.step(eps, t, xo)
        x, hist = _emulate_kernel(s._table[i], x, eps, hist if s._uses_history else None)
        torch.testing.assert_close(x, xo, rtol=2e-5, atol=2e-5)


class _KernelState:
    """fp32 emulation of cfg_step_kernel (csrc/bandwidth.cu) including the noise, eps-history and saved-sample terms."""

    def __init__(self, x):
        self.hist = torch.zeros_like(x)
        self.eh = [torch.zeros_like(x) for _ in range(3)]
        self.saved = torch.zeros_like(x)

    def step(self, row, x, eps, noise, sched):
        g = lambda c: row[c]
        flags = int(row[N.ST_FLAGS])
        xs = self.saved if (sched._uses_eps_history and flags & N.STF_USE_SAVED_X) else x
        x0 = (xs - g(N.ST_K1) * eps) * g(N.ST_K2)
        xn = g(N.ST_CX) * xs + g(N.ST_CE) * eps + g(N.ST_C0) * x0
        if flags & N.STF_UNIPC:
            x0 = (x - g(N.ST_K1) * eps) * g(N.ST_K2)
            m1, m2 = self.hist, self.eh[0]
            xc = x
            if flags & N.STF_UNIPC_CORRECT:
                xc = g(N.ST_H1) * self.saved + g(N.ST_H2) * m1 + g(N.ST_H3) * (m2 - m1) + g(N.ST_CE) * (x0 - m1)
            xn = g(N.ST_CX) * xc + g(N.ST_C0) * x0 + g(N.ST_C1) * (x0 - m1)
            self.saved, self.eh[0], self.hist = xc, m1, x0
            return xn
        if sched._uses_history:
            if g(N.ST_C1) != 0:
                xn = xn + g(N.ST_C1) * (x0 - self.hist)
            self.hist = x0
        if sched._stochastic and g(N.ST_CN) != 0:
            xn = xn + g(N.ST_CN) * noise
        if sched._uses_eps_history:
            xn = xn + g(N.ST_H1) * self.eh[0] + g(N.ST_H2) * self.eh[1] + g(N.ST_H3) * self.eh[2]
            if flags & N.STF_PUSH_EPS:
                self.eh = [eps, self.eh[0], self.eh[1]]
            if flags & N.STF_SAVE_X:
                self.saved = x
        return xn


@pytest.mark.parametrize("name,n", [("euler_a", 20), ("euler_a", 1), ("DPM++ 2M SDE Karras", 20), ("DPM++ 2M SDE Karras", 2),
                                    ("PNDM", 20), ("PNDM", 50), ("PNDM", 3), ("euler", 7), ("DPM++ 2M", 9),
                                    ("uni_pc", 20), ("uni_pc", 50), ("uni_pc", 2), ("uni_pc", 1), ("uni_pc", 3)])
def test_step_table_next_row_schedulers(name, n):
    """the table rows of euler_a / DPM++ 2M SDE / PNDM drive the kernel's linear form to the oracle's
    trajectory (PNDM: N + 1 evaluations, second timestep twice, averaged redo of the first step)."""
    s = make_scheduler(name, B200EulerDiscreteScheduler(**CFG).config)
    s.set_timesteps(n)
    o = OS.make_scheduler(name, CFG)
    o.set_timesteps(n)
    assert torch.equal(s.timesteps.cpu(), o.timesteps) and len(s._table) == len(o.timesteps)
    g = torch.Generator().manual_seed(1)
    x = torch.randn(2, 4, 8, 8, generator=g) * float(o.init_noise_sigma)
    xo, ks = x.clone(), _KernelState(x)
    for i, t in enumerate(o.timesteps):
        eps = torch.randn(2, 4, 8, 8, generator=g)
        noise = torch.randn(2, 4, 8, 8, generator=g)
        assert s._table[i, N.ST_TIMESTEP].item() == float(t)
        xo = o.step(eps, t, xo, noise) if name in OS.STOCHASTIC else o.step(eps, t, xo)
        x = ks.step(s._table[i], x, eps, noise, s)
        torch.testing.assert_close(x, xo, rtol=5e-5, atol=5e-5)


def test_pndm_rows_follow_the_call_counter():
    """img2img starts mid-schedule: PNDM's special first/second step follow its call counter, not the table position."""
    s = make_scheduler("PNDM", B200EulerDiscreteScheduler(**CFG).config)
    s.set_timesteps(20)
    rows = s.rows_for(8, 5)
    assert rows.shape[0] == 5 and int(rows[0, N.ST_FLAGS]) == (N.STF_PUSH_EPS | N.STF_SAVE_X)
    assert int(rows[1, N.ST_FLAGS]) == N.STF_USE_SAVED_X and rows[0, N.ST_TIMESTEP] == float(s.timesteps[8])
    assert s.timesteps.tolist()[:4] == [951, 901, 901, 851] and len(s.timesteps) == 21


def test_from_config_uses_target_class_defaults():
    e = B200EulerDiscreteScheduler(**CFG)                       # SD1.5's PNDM-style config: no timestep_spacing key
    assert e.config.timestep_spacing == "linspace"
    d = B200DDIMScheduler.from_config(e.config)
    assert d.config.timestep_spacing == "leading" and d.config.steps_offset == 1 and d.config.clip_sample is False
    m = make_scheduler("DPM++ 2M Karras", d.config)
    assert isinstance(m, B200DPMSolverMultistepScheduler) and m.config.use_karras_sigmas and m.config.beta_schedule == "scaled_linear"
    with pytest.raises(ValueError, match="Unknown scheduler name"):
        make_scheduler("nope", e.config)
    for name in ("euler", "euler_a", "DDIM", "DPM++ 2M", "DPM++ 2M Karras", "DPM++ 2M SDE Karras", "PNDM", "uni_pc"):
        make_scheduler(name, e.config)                           # the whole name table of the reference (:205-222)


def test_geglu_interleave_and_conv_packing():
    from stablediffusion_b200.packing import conv_weight_to_gemm, geglu_block_n, geglu_interleave

    C = 320
    w = torch.arange(8 * C, dtype=torch.float32)[:, None].repeat(1, 2)
    bn = geglu_block_n(8 * C)
    p = geglu_interleave(w, bn)
    half = bn // 2
    for t in range(8 * C // bn):
        assert torch.equal(p[t * bn:t * bn + half, 0], torch.arange(t * half, (t + 1) * half, dtype=torch.float32))
        assert torch.equal(p[t * bn + half:(t + 1) * bn, 0], 4 * C + torch.arange(t * half, (t + 1) * half, dtype=torch.float32))
    w4 = torch.randn(6, 5, 3, 3)
    g = conv_weight_to_gemm(w4)
    assert g.shape == (6, 45) and torch.equal(g[:, (1 * 3 + 2) * 5 + 4], w4[:, 4, 1, 2])   # [co][kh][kw][ci]


def _fake_lora(paths, r=4, dims=None, seed=0, fmt="diffusers", alpha=None):
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for p in paths:
        o, i = dims[p]
        A, B = torch.randn(r, i, generator=g) / r, torch.randn(o, r, generator=g) * 0.02
        if fmt == "diffusers":     # train_lora_pipeline.py:497-528 -> save_lora_weights
            sd[f"unet.{p}.lora.down.weight"], sd[f"unet.{p}.lora.up.weight"] = A, B
        elif fmt == "peft":
            sd[f"unet.{p}.lora_A.weight"], sd[f"unet.{p}.lora_B.weight"] = A, B
        else:                      # A1111 / kohya
            k = "lora_unet_" + p.replace(".", "_")
            sd[f"{k}.lora_down.weight"], sd[f"{k}.lora_up.weight"] = A, B
            sd[f"{k}.alpha"] = torch.tensor(float(alpha))
    return sd


@pytest.mark.parametrize("fmt,alpha", [("diffusers", None), ("peft", None), ("kohya", 2.0)])
def test_lora_merge_equals_unmerged_forward(fmt, alpha):
    """merged-weights forward == PEFT-semantics unmerged forward, fp32, rel <= 1e-5 (SURVEY §4 test plan)."""
    from oracle.lora import attach_unmerged
    from stablediffusion_b200.lora import merged_state_dict, parse_lora_state_dict
    from stablediffusion_b200.unet import unet_param_shapes
    from tests.golden.make_golden import SMALL_UNET, build_small_unet

    shapes = unet_param_shapes(unet_config(**SMALL_UNET))
    targets = [k[:-7] for k in shapes if k.endswith((".to_q.weight", ".to_k.weight", ".to_v.weight", ".to_out.0.weight"))]
    dims = {p: tuple(shapes[p + ".weight"]) for p in targets}
    modules = sorted({k.rsplit(".", 1)[0] for k in shapes if k.endswith(".weight")})
    a1 = parse_lora_state_dict(_fake_lora(targets, 4, dims, 1, fmt, alpha), unet_modules=modules)["unet"]
    a2 = parse_lora_state_dict(_fake_lora(targets[:6], 2, dims, 2, "diffusers"), unet_modules=modules)["unet"]
    assert len(a1) == len(targets) == 6 * 8 // 2 * 1 or len(a1) == len(targets)
    adapters, weights = {"style": a1, "char": a2}, {"style": 0.8, "char": 0.5}
    base = build_small_unet()
    merged = build_small_unet()
    merged.load_state_dict(merged_state_dict(base.state_dict(), adapters, weights), strict=True)
    attach_unmerged(base, adapters, weights)
    g = torch.Generator().manual_seed(3)
    x, c = torch.randn(1, 4, 16, 16, generator=g), torch.randn(1, 77, 64, generator=g)
    with torch.no_grad():
        a, b = base(x, torch.tensor(500.0), c), merged(x, torch.tensor(500.0), c)
    assert ((a - b).norm() / a.norm()).item() <= 1e-5
    plain = build_small_unet()
    with torch.no_grad():
        assert ((plain(x, torch.tensor(500.0), c) - a).norm() / a.norm()).item() > 1e-4   # the adapters do something


def test_invalid_lora_checkpoint_raises():
    from stablediffusion_b200.lora import parse_lora_state_dict

    with pytest.raises(ValueError, match="Invalid LoRA checkpoint"):
        parse_lora_state_dict({"unet.conv_in.weight": torch.zeros(1)})


def test_img2img_timestep_slice():
    """reference :722-761: N=50, strength 0.6 -> 30 steps starting at index 20."""
    from stablediffusion_b200.pipelines.sd_unified_pipeline import StableDiffusionUnifiedPipeline

    class M:
        pass

    p = StableDiffusionUnifiedPipeline(device="cuda")
    p.model = M()
    p.model.scheduler = B200EulerDiscreteScheduler(**CFG)
    p.model.scheduler.set_timesteps(50)
    ts, n = p.get_timesteps(50, 0.6)
    assert n == 30 and len(ts) == 30 and torch.equal(ts, p.model.scheduler.timesteps[20:])
    ts, n = p.get_timesteps(50, 1.0)
    assert n == 50
    ts, n = p.get_timesteps(50, 0.0)
    assert n == 0 and len(ts) == 0


def test_param_shapes_match_oracle_strictly():
    from oracle.unet import SD15_UNET, SDXL_UNET, UNet2DConditionOracle
    from stablediffusion_b200.unet import unet_param_shapes

    for mt, oc in (("sd15", SD15_UNET), ("sdxl", SDXL_UNET)):
        with torch.device("meta"):
            osd = {k: tuple(v.shape) for k, v in UNet2DConditionOracle(**oc).state_dict().items()}
        assert osd == {k: tuple(v) for k, v in unet_param_shapes(unet_config(mt)).items()}


# ---------------------------------------------------------------------------------------------------
# round-2 advisor findings
# ---------------------------------------------------------------------------------------------------
ALL_NAMES = ("euler", "euler_a", "DDIM", "DPM++ 2M", "DPM++ 2M Karras", "DPM++ 2M SDE Karras", "PNDM", "uni_pc")


@pytest.mark.parametrize("model_type", ["sd15", "sdxl"])
@pytest.mark.parametrize("name", ALL_NAMES)
def test_every_scheduler_name_builds_from_both_model_configs(name, model_type):
    """The reference's set_scheduler (models/stable_diffusion.py:199-227) accepts every name on SD1.5 AND SDXL; the SDXL
    scheduler config carries timestep_spacing="leading" / steps_offset=1 into every class through from_config."""
    base = B200EulerDiscreteScheduler(**scheduler_config(model_type))
    s = make_scheduler(name, base.config)
    s.set_timesteps(30)
    o = OS.make_scheduler(name, dict(OS.SDXL_SCHEDULER_CONFIG if model_type == "sdxl" else OS.SD15_SCHEDULER_CONFIG))
    o.set_timesteps(30)
    assert s.timesteps.dtype == o.timesteps.dtype and torch.equal(s.timesteps.cpu(), o.timesteps)
    if model_type == "sdxl" and name in ("DPM++ 2M", "uni_pc", "DPM++ 2M SDE Karras"):
        # leading spacing of the multistep solvers: arange(0, n+1) * (1000 // (n+1)), reversed, last dropped, + steps_offset
        assert s.timesteps.tolist()[:3] == [961, 929, 897] and s.timesteps.tolist()[-1] == 33


@pytest.mark.parametrize("name", ["DPM++ 2M", "uni_pc", "DPM++ 2M SDE Karras", "DPM++ 2M Karras"])
def test_multistep_solvers_on_the_sdxl_config_follow_the_oracle(name):
    s = make_scheduler(name, B200EulerDiscreteScheduler(**scheduler_config("sdxl")).config)
    s.set_timesteps(12)
    o = OS.make_scheduler(name, dict(OS.SDXL_SCHEDULER_CONFIG))
    o.set_timesteps(12)
    g = torch.Generator().manual_seed(3)
    x = torch.randn(1, 4, 8, 8, generator=g)
    xo, ks = x.clone(), _KernelState(x)
    for i, t in enumerate(o.timesteps):
        eps, noise = torch.randn(1, 4, 8, 8, generator=g), torch.randn(1, 4, 8, 8, generator=g)
        xo = o.step(eps, t, xo, noise) if name in OS.STOCHASTIC else o.step(eps, t, xo)
        x = ks.step(s._table[i], x, eps, noise, s)
        torch.testing.assert_close(x, xo, rtol=5e-5, atol=5e-5)


def test_unfused_multistep_first_call_is_first_order_wherever_it_starts():
    """img2img hands scheduler.step a timestep from the middle of the schedule: diffusers runs that first call
    first-order (no history yet), so the un-fused table must carry C1 = 0 on ITS first row, not on row 0 only."""
    s = make_scheduler("DPM++ 2M", B200EulerDiscreteScheduler(**CFG).config)
    s.set_timesteps(20)
    assert s._table[8, N.ST_C1] != 0
    s._init_step_index(s.timesteps[8])
    tab = s._unfused_table("cpu")
    assert tab[8, N.ST_C1] == 0 and tab[9, N.ST_C1] == s._table[9, N.ST_C1] != 0
    o = OS.make_scheduler("DPM++ 2M", CFG)
    o.set_timesteps(20)
    g = torch.Generator().manual_seed(5)
    x = torch.randn(1, 4, 8, 8, generator=g)
    xo, ks = x.clone(), _KernelState(x)
    for i in range(8, 20):
        eps = torch.randn(1, 4, 8, 8, generator=g)
        xo = o.step(eps, o.timesteps[i], xo)
        x = ks.step(tab[i], x, eps, None, s)
        torch.testing.assert_close(x, xo, rtol=5e-5, atol=5e-5)


def test_lora_with_unresolvable_modules_raises_instead_of_loading_as_a_noop():
    from stablediffusion_b200.lora import parse_lora_state_dict

    mods = ["down_blocks.0.attentions.0.transformer_blocks.0.attn1.to_q"]
    good = {"lora_unet_down_blocks_0_attentions_0_transformer_blocks_0_attn1_to_q.lora_down.weight": torch.zeros(4, 8),
            "lora_unet_down_blocks_0_attentions_0_transformer_blocks_0_attn1_to_q.lora_up.weight": torch.zeros(8, 4)}
    assert list(parse_lora_state_dict(good, unet_modules=mods)["unet"]) == mods
    sgm = {"lora_unet_input_blocks_1_1_transformer_blocks_0_attn1_to_q.lora_down.weight": torch.zeros(4, 8),
           "lora_unet_input_blocks_1_1_transformer_blocks_0_attn1_to_q.lora_up.weight": torch.zeros(8, 4)}
    with pytest.raises(ValueError, match="does not have"):
        parse_lora_state_dict(sgm, unet_modules=mods)


def test_text_encoder_lora_kohya_names_and_exact_restore():
    """kohya text-encoder keys resolve through the flattened module names; load / set / delete cycles re-merge from an fp32
    pristine copy, so the encoder's weights come back bit for bit."""
    from stablediffusion_b200.models.stable_diffusion import SDModelWrapper

    class Enc(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.text_model = torch.nn.Module()
            self.text_model.q_proj = torch.nn.Linear(8, 8, bias=False).to(torch.bfloat16)

    enc = Enc()
    w0 = enc.text_model.q_proj.weight.data.clone()
    w = SDModelWrapper.__new__(SDModelWrapper)
    layers = {"lora_te_text_model_q_proj": dict(down=torch.randn(2, 8), up=torch.randn(8, 2), alpha=2.0)}
    w._merge_text_encoder_lora(enc, layers, "a")
    for wt in (0.7, 0.3, 1.0, 0.0):
        w._apply_text_encoder_weights(enc, ["a"], [wt])
    assert torch.equal(enc.text_model.q_proj.weight.data, w0)
    w._apply_text_encoder_weights(enc, ["a"], [1.0])
    assert not torch.equal(enc.text_model.q_proj.weight.data, w0)
    with pytest.raises(ValueError, match="do not exist"):
        w._merge_text_encoder_lora(enc, {"lora_te_text_model_nope": layers["lora_te_text_model_q_proj"]}, "b")


def test_cross_attention_scale_is_not_silently_ignored():
    from stablediffusion_b200.pipelines.sd_unified_pipeline import StableDiffusionUnifiedPipeline

    pipe = StableDiffusionUnifiedPipeline(do_cfg=True, device="cuda")
    with pytest.raises(ValueError, match="set_adapters"):
        pipe(object(), prompt="x", cross_attention_kwargs={"scale": 0.5})


def test_unet_pack_version_is_monotonic():
    from stablediffusion_b200.unet import B200UNet

    u = B200UNet(unet_config(), None, device="cpu")
    assert u.pack_version == 0


def test_upsample_conv_phase_weights_reproduce_conv_of_upsampled_input():
    """Upsample2D folded into its conv (packing.upsample_conv_phase_weights): four 2x2-tap phase convs over the low-resolution
    input, each filling one lattice of the output, equal conv3x3(nearest_upsample_2x(x)) exactly (fp64), border included."""
    import torch.nn.functional as F

    from stablediffusion_b200.packing import upsample_conv_phase_taps, upsample_conv_phase_weights

    g = torch.Generator().manual_seed(0)
    NB, H, W, Ci, Co = 2, 5, 7, 8, 6
    x = torch.randn(NB, Ci, H, W, generator=g, dtype=torch.float64)
    w = torch.randn(Co, Ci, 3, 3, generator=g, dtype=torch.float64)
    ref = F.conv2d(F.interpolate(x, scale_factor=2.0, mode="nearest"), w, padding=1)
    wp = upsample_conv_phase_weights(w)
    assert wp.shape == (4, Co, 4 * Ci)
    out = torch.zeros_like(ref)
    xp = F.pad(x, (1, 1, 1, 1))
    for py in (0, 1):
        for px in (0, 1):
            acc = torch.zeros(NB, Co, H, W, dtype=torch.float64)
            for t, (ph, dh, dw) in enumerate(upsample_conv_phase_taps(py, px)):
                assert ph == 0 and dh in (-1, 0, 1) and dw in (-1, 0, 1)
                acc += torch.einsum("oc,nchw->nohw", wp[py * 2 + px][:, t * Ci:(t + 1) * Ci], xp[:, :, 1 + dh:1 + dh + H, 1 + dw:1 + dw + W])
            out[:, :, py::2, px::2] = acc
    assert (out - ref).abs().max().item() < 1e-12


def test_padding_mask_crop_region_and_fill_resize():
    """`padding_mask_crop` (reference :270-285): the crop region holds the padded mask box, has the processing aspect ratio,
    stays inside the image; image and mask are cropped to it and brought to (width, height)."""
    import numpy as np
    from PIL import Image

    from stablediffusion_b200.models.stable_diffusion import _ImageProcessor

    proc = _ImageProcessor(8)
    m = np.zeros((300, 400), dtype=np.uint8)
    m[100:160, 250:330] = 255
    mask = Image.fromarray(m)
    x1, y1, x2, y2 = proc.get_crop_region(mask, 512, 512, pad=16)
    assert x1 <= 250 - 16 and x2 >= 330 + 16 and y1 <= 100 - 16 and y2 >= 160 + 16
    assert 0 <= x1 < x2 <= 400 and 0 <= y1 < y2 <= 300
    assert abs((x2 - x1) / (y2 - y1) - 1.0) < 0.05
    # a box at the border is shifted back inside, not clipped out of ratio
    m2 = np.zeros((300, 400), dtype=np.uint8)
    m2[0:40, 380:400] = 255
    bx = proc.get_crop_region(Image.fromarray(m2), 512, 256, pad=8)
    assert bx[0] >= 0 and bx[2] <= 400 and bx[1] >= 0 and bx[3] <= 300 and abs((bx[2] - bx[0]) / (bx[3] - bx[1]) - 2.0) < 0.1
    img = Image.fromarray((np.random.default_rng(0).random((300, 400, 3)) * 255).astype(np.uint8))
    t = proc.preprocess(img, height=512, width=512, crops_coords=(x1, y1, x2, y2), resize_mode="fill")
    assert t.shape == (1, 3, 512, 512) and -1.0 <= float(t.min()) and float(t.max()) <= 1.0
    mk = _ImageProcessor(8, do_normalize=False).preprocess_mask(mask, height=512, width=512, crops_coords=(x1, y1, x2, y2), resize_mode="fill")
    assert mk.shape == (1, 1, 512, 512) and set(mk.unique().tolist()) <= {0.0, 1.0}
    assert 0.05 < mk.mean().item() < 0.6        # the masked box now fills a good part of the frame
    with pytest.raises(ValueError, match="PIL"):
        proc.preprocess(torch.zeros(1, 3, 64, 64), crops_coords=(0, 0, 8, 8))
