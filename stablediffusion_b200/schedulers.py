"""B200 schedulers: the replacement for `model.scheduler` on the reference's hot path
(scale_model_input pipelines/sd_unified_pipeline.py:472, step :489, add_noise :502,841,890, set_timesteps via
retrieve_timesteps :61-95; classes selected in models/stable_diffusion.py:135-138,199-227).

Host side: the timestep / sigma tables are built with the same numpy expressions diffusers 0.27.2 uses, so
integer timesteps are exact and float tables are bit-identical.  Every per-step scalar is
folded into one row of a device "step table"
    [t, in_scale, K1, K2, CX, CE, C0, C1]:   x0 = (x - K1*eps)*K2 ;  x' = CX*x + CE*eps + C0*x0 + C1*(x0 - x0_prev)
which the fused CFG+step kernel (libb200sd OP_CFG_STEP) indexes with a device-side step counter — so one
captured CUDA graph serves every sampler step.  The coefficient scalars are computed in torch fp32 in
diffusers' operation order.

Device side: `scale_model_input`, `step` and `add_noise` called the reference's way (un-fused, one call per
step) run the same kernels through a tiny per-shape op program.
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import numpy as np
import torch

from . import _native as N
from . import lowering as L
from .config import Config, scheduler_config


def _betas(cfg) -> torch.Tensor:
    n = cfg["num_train_timesteps"]
    if cfg.get("trained_betas") is not None:
        return torch.tensor(cfg["trained_betas"], dtype=torch.float32)
    sched = cfg.get("beta_schedule", "linear")
    if sched == "scaled_linear":
        return torch.linspace(cfg["beta_start"] ** 0.5, cfg["beta_end"] ** 0.5, n, dtype=torch.float32) ** 2
    if sched == "linear":
        return torch.linspace(cfg["beta_start"], cfg["beta_end"], n, dtype=torch.float32)
    raise ValueError(f"{sched} is not implemented")


class _SchedulerBase:
    order = 1
    _class_defaults: dict = {}

    def __init__(self, **config):
        c = scheduler_config()
        c.update(self._class_defaults)
        explicit = {k: v for k, v in config.items() if k != "_use_default_values"}
        c.update(explicit)
        # like diffusers' ConfigMixin: remember which entries are class defaults so that from_config() on ANOTHER
        # scheduler class lets that class's own defaults win for them
        c["_use_default_values"] = sorted(set(c) - set(explicit))
        self.config = c
        self.betas = _betas(c)
        self.alphas = 1.0 - self.betas
        self.alphas_cumprod = torch.cumprod(self.alphas, dim=0)
        self.timesteps: Optional[torch.Tensor] = None   # on the device given to set_timesteps
        self._timesteps_host: Optional[np.ndarray] = None
        self.num_inference_steps: Optional[int] = None
        self._step_index: Optional[int] = None
        self._table: Optional[torch.Tensor] = None       # host fp32 [n, STEP_ROW]
        self._dev: Dict[str, torch.Tensor] = {}
        self._programs: Dict[tuple, dict] = {}
        self._ew_programs: Dict[tuple, dict] = {}        # shape-keyed AXPBY programs (independent of the timestep table)
        self._coef_dev: Dict[tuple, torch.Tensor] = {}   # device copies of the (a, b) pairs they are run with
        self._start = 0                                  # schedule position of the first un-fused step() call
        self.device = torch.device("cpu")

    @classmethod
    def from_config(cls, config, **kwargs):
        """`XScheduler.from_config(self.scheduler.config)` (models/stable_diffusion.py:203-222): keys the class
        does not know are ignored, class defaults fill the rest."""
        dropped = set(config.get("_use_default_values", ())) | {"_use_default_values"}
        c = {k: v for k, v in dict(config).items() if k not in dropped}
        c.update(kwargs)
        return cls(**c)

    # ---- step-table plumbing -----------------------------------------------------------------------
    def _finish_set_timesteps(self, device):
        self.device = torch.device(device) if device is not None else torch.device("cpu")
        self._step_index = None
        self._dev.clear()
        self._programs.clear()
        self._table = self._build_table()

    def step_table(self, device, start_index: int = 0) -> torch.Tensor:
        """Device copy of the table rows [start_index:], plus the step counter (int32[1])."""
        key = f"table{start_index}@{device}"
        if key not in self._dev:
            self._dev[key] = self._table[start_index:].contiguous().to(device)
        return self._dev[key]

    @property
    def step_index(self):
        return self._step_index

    def index_for_timestep(self, timestep) -> int:
        t = float(timestep)
        idx = np.nonzero(self._timesteps_host.astype(np.float64) == t)[0]
        if len(idx) == 0:
            raise ValueError(f"timestep {t} is not in the schedule")
        return int(idx[1] if len(idx) > 1 else idx[0])

    def _init_step_index(self, timestep):
        if self._step_index is None:
            self._step_index = self._start = self.index_for_timestep(timestep)

    # ---- un-fused device calls (the reference's call pattern) ----------------------------------------
    def _elementwise(self, x, a, y=None, bcoef=0.0):
        """out = a*x + b*y via the library's AXPBY kernel: ONE planned op per (shape, dtype), its coefficients read from
        device memory, enqueued on the caller's stream without any host synchronisation."""
        if x.device.type != "cuda":
            raise RuntimeError("B200 schedulers operate on CUDA tensors only (no CPU path)")
        key = ("axpby", tuple(x.shape), x.dtype, y is not None, x.device)
        st = self._ew_programs.get(key)
        if st is None:
            b = L.Builder(x.device, workspace_bytes=16)
            st = dict(b=b, x=b.empty(x.shape, x.dtype), y=b.empty(x.shape, x.dtype) if y is not None else None,
                      coef=b.zeros([2], torch.float32), out=b.empty(x.shape, x.dtype))
            b.axpby(st["x"], 1.0, y=st["y"], b=0.0, out=st["out"], coef=st["coef"])
            self._ew_programs[key] = st
        st["x"].copy_(x)
        if y is not None:
            st["y"].copy_(y)
        # the coefficient pair lives on the device: uploaded once per distinct (a, b) — i.e. once per sampler step of the
        # first trajectory of a schedule — and moved into the op's slot by a device-to-device copy afterwards
        ck = (float(a), float(bcoef), x.device)
        cd = self._coef_dev.get(ck)
        if cd is None:
            if len(self._coef_dev) > 4096:
                self._coef_dev.clear()
            cd = self._coef_dev[ck] = torch.tensor([float(a), float(bcoef)], dtype=torch.float32, device=x.device)
        st["coef"].copy_(cd)
        st["b"].prog.run(torch.cuda.current_stream(x.device).cuda_stream)
        return st["out"].clone()

    def scale_model_input(self, sample, timestep=None):
        return sample

    def step(self, model_output, timestep, sample, return_dict=False, generator=None, **unused):
        """x_{t-1} from (already CFG-combined) eps — the un-fused call at pipeline :489.  Stochastic samplers draw
        their noise here, like diffusers, from `generator` (the reference passes none: the device's global RNG)."""
        self._init_step_index(timestep)
        dev = sample.device
        if dev.type != "cuda":
            raise RuntimeError("B200 schedulers operate on CUDA tensors only (no CPU path)")
        B, C, h, w = sample.shape
        key = ("step", self._start, B, C, h, w, sample.dtype, model_output.dtype)
        st = self._programs.get(key)
        if st is None:
            b = L.Builder(dev, workspace_bytes=16)
            st = dict(b=b)
            st["eps_nchw"] = b.empty([B, C, h, w], model_output.dtype)
            st["eps"] = b.empty([B, h * w, 8])
            b.prep_latents(st["eps_nchw"], st["eps"], rep=1, fixed_scale=1.0)  # NCHW -> token-major bf16
            st["x"] = b.empty([B, C, h, w], torch.float32)
            st["counter"] = b.zeros([1], torch.int32)
            st["hist"] = b.zeros([B, C, h * w], torch.float32) if self._uses_history else None
            st["eps_hist"] = b.zeros([3, B, C, h * w], torch.float32) if self._uses_eps_history else None
            st["x_saved"] = b.zeros([B, C, h * w], torch.float32) if self._uses_eps_history else None
            st["noise"] = b.zeros([len(self._table), B, C, h * w], torch.float32) if self._stochastic else None
            b.cfg_step(st["eps"], st["x"], st["x"], self._unfused_table(dev), st["counter"], 1.0, False,
                       len(self._table), x0_prev=st["hist"], noise=st["noise"], eps_hist=st["eps_hist"], x_saved=st["x_saved"])
            self._programs[key] = st
        cidx = self._unfused_counter()
        st["eps_nchw"].copy_(model_output)
        st["x"].copy_(sample)
        st["counter"].fill_(cidx)
        if self._stochastic:
            st["noise"][cidx].copy_(torch.randn((B, C, h, w), generator=generator, device=dev, dtype=torch.float32).view(B, C, h * w))
        st["b"].prog.run(torch.cuda.current_stream(dev).cuda_stream)
        self._step_index += 1
        # diffusers casts the result to the model-output dtype; we return the sample's dtype when
        # it is wider (fp32 latents are strictly more accurate) and the model dtype otherwise.
        out_dtype = sample.dtype if sample.dtype == torch.float32 else model_output.dtype
        prev = st["x"].to(out_dtype)
        return (prev,)

    def _unfused_table(self, dev):
        """Table of the un-fused calls.  A multistep solver has no history on its FIRST call wherever that lands in the
        schedule (img2img starts mid-table; diffusers runs that step first-order: lower_order_nums == 0), so the row of
        the first executed position gets C1 = 0 — the same thing DenoiseEngine.load does for the fused loop."""
        key = f"unfused{self._start}@{dev}"
        if key not in self._dev:
            tab = self._table.clone()
            tab[self._start, N.ST_C1] = 0.0
            self._dev[key] = tab.to(dev)
        return self._dev[key]

    def _unfused_counter(self) -> int:
        return self._step_index

    def rows_for(self, start_index: int, n: int) -> torch.Tensor:
        """Host table rows of the n steps executed from `start_index` (img2img starts mid-schedule)."""
        return self._table[start_index:start_index + n]

    def noise_coeffs(self, timestep):
        """(a, s) of `add_noise`: x_t = a * x_0 + s * noise at `timestep` (also the inpaint blend's re-noising, :499-504)."""
        raise NotImplementedError

    def add_noise(self, original_samples, noise, timesteps):
        ts = set(float(t) for t in timesteps.reshape(-1).tolist())
        if len(ts) != 1:
            raise ValueError("add_noise: per-sample timesteps must be equal on this path")
        a, s_ = self.noise_coeffs(ts.pop())
        return self._elementwise(original_samples, float(a), noise, float(s_))

    _uses_eps_history = False   # PNDM / PLMS: eps history + remembered sample
    _stochastic = False         # the update has a noise term (ancestral / SDE samplers)
    _uses_history = False


class B200EulerDiscreteScheduler(_SchedulerBase):
    """EulerDiscreteScheduler (name `euler`, the default: models/stable_diffusion.py:100,135)."""
    _class_defaults = dict(timestep_spacing="linspace", interpolation_type="linear", use_karras_sigmas=False,
                           timestep_type="discrete", steps_offset=0)

    def __init__(self, **config):
        super().__init__(**config)
        sig = ((1 - self.alphas_cumprod) / self.alphas_cumprod) ** 0.5
        self.sigmas = torch.cat([sig.flip(0), torch.zeros(1)])

    @property
    def init_noise_sigma(self):
        m = self.sigmas.max()
        if self.config["timestep_spacing"] in ("linspace", "trailing"):
            return m
        return (m ** 2 + 1) ** 0.5

    def set_timesteps(self, num_inference_steps, device=None):
        c = self.config
        ntrain = c["num_train_timesteps"]
        n = self.num_inference_steps = int(num_inference_steps)
        sp = c["timestep_spacing"]
        if sp == "linspace":
            ts = np.linspace(0, ntrain - 1, n, dtype=np.float32)[::-1].copy()
        elif sp == "leading":
            ts = (np.arange(0, n) * (ntrain // n)).round()[::-1].copy().astype(np.float32)
            ts += c["steps_offset"]
        elif sp == "trailing":
            ts = (np.arange(ntrain, 0, -ntrain / n)).round().copy().astype(np.float32)
            ts -= 1
        else:
            raise ValueError(f"{sp} is not supported")
        sig = (((1 - self.alphas_cumprod) / self.alphas_cumprod) ** 0.5).numpy()
        sig = np.interp(ts, np.arange(0, len(sig)), sig)
        sig = torch.from_numpy(sig).to(dtype=torch.float32)
        self._timesteps_host = ts.astype(np.float32)
        self.sigmas = torch.cat([sig, torch.zeros(1)])
        self._finish_set_timesteps(device)
        self.timesteps = torch.from_numpy(self._timesteps_host).to(self.device)

    def _build_table(self):
        n = self.num_inference_steps
        tab = torch.zeros(n, N.STEP_ROW, dtype=torch.float32)
        for i in range(n):
            s, s_next = self.sigmas[i], self.sigmas[i + 1]
            tab[i, N.ST_TIMESTEP] = float(self._timesteps_host[i])
            tab[i, N.ST_IN_SCALE] = 1.0 / ((s ** 2 + 1) ** 0.5)
            # x0 = x - s*eps ; d = (x - x0)/s ; x' = x + d*(s_next - s)  ==  x + eps*dt up to fp32 rounding
            tab[i, N.ST_K1], tab[i, N.ST_K2] = s, 1.0
            tab[i, N.ST_CX], tab[i, N.ST_CE] = 1.0, s_next - s
        return tab

    def scale_model_input(self, sample, timestep=None):
        self._init_step_index(timestep)
        s = self.sigmas[self._step_index]
        return self._elementwise(sample, float(1.0 / ((s ** 2 + 1) ** 0.5)))

    def noise_coeffs(self, timestep):
        return 1.0, float(self.sigmas[self.index_for_timestep(timestep)])


class B200DDIMScheduler(_SchedulerBase):
    """DDIMScheduler (name `DDIM`, BASELINE config 2), eta = 0."""
    _class_defaults = dict(timestep_spacing="leading", clip_sample=True, set_alpha_to_one=True, steps_offset=0,
                           thresholding=False)
    init_noise_sigma = 1.0

    def __init__(self, **config):
        super().__init__(**config)
        if self.config["clip_sample"] or self.config.get("thresholding"):
            raise ValueError("clip_sample / thresholding DDIM is not part of the SD hot path (SD configs set clip_sample=False)")
        self.final_alpha_cumprod = torch.tensor(1.0) if self.config["set_alpha_to_one"] else self.alphas_cumprod[0]

    def set_timesteps(self, num_inference_steps, device=None):
        c = self.config
        ntrain = c["num_train_timesteps"]
        n = self.num_inference_steps = int(num_inference_steps)
        sp = c["timestep_spacing"]
        if sp == "leading":
            ts = (np.arange(0, n) * (ntrain // n)).round()[::-1].copy().astype(np.int64)
            ts += c["steps_offset"]
        elif sp == "linspace":
            ts = np.linspace(0, ntrain - 1, n).round()[::-1].copy().astype(np.int64)
        elif sp == "trailing":
            ts = np.round(np.arange(ntrain, 0, -ntrain / n)).astype(np.int64)
            ts -= 1
        else:
            raise ValueError(f"{sp} is not supported")
        self._timesteps_host = ts
        self._finish_set_timesteps(device)
        self.timesteps = torch.from_numpy(ts).to(self.device)

    def _build_table(self):
        n = self.num_inference_steps
        ntrain = self.config["num_train_timesteps"]
        tab = torch.zeros(n, N.STEP_ROW, dtype=torch.float32)
        for i in range(n):
            t = int(self._timesteps_host[i])
            prev_t = t - ntrain // n
            a_t = self.alphas_cumprod[t]
            a_p = self.alphas_cumprod[prev_t] if prev_t >= 0 else self.final_alpha_cumprod
            tab[i, N.ST_TIMESTEP] = float(t)
            tab[i, N.ST_IN_SCALE] = 1.0
            tab[i, N.ST_K1] = (1 - a_t) ** 0.5          # x0 = (x - sqrt(1-a_t) eps) / sqrt(a_t)
            tab[i, N.ST_K2] = 1.0 / (a_t ** 0.5)
            tab[i, N.ST_CE] = (1 - a_p) ** 0.5          # x' = sqrt(a_p) x0 + sqrt(1-a_p) eps
            tab[i, N.ST_C0] = a_p ** 0.5
        return tab

    def noise_coeffs(self, timestep):
        a = self.alphas_cumprod[int(timestep)]
        return float(a ** 0.5), float((1 - a) ** 0.5)


class B200DPMSolverMultistepScheduler(_SchedulerBase):
    """DPMSolverMultistepScheduler as `DPM++ 2M` / `DPM++ 2M Karras` (models/stable_diffusion.py:211-214):
    dpmsolver++, order 2, midpoint, final sigma zero."""
    _class_defaults = dict(timestep_spacing="linspace", algorithm_type="dpmsolver++", solver_order=2, solver_type="midpoint",
                           lower_order_final=True, use_karras_sigmas=False, final_sigmas_type="zero", steps_offset=0)
    init_noise_sigma = 1.0
    _uses_history = True

    def __init__(self, **config):
        super().__init__(**config)
        c = self.config
        if c["algorithm_type"] != "dpmsolver++" or c["solver_order"] != 2 or c["solver_type"] != "midpoint":
            raise ValueError("only dpmsolver++ / order 2 / midpoint (DPM++ 2M) is on the hot path; the SDE variant is a 'next' row")
        if c["timestep_spacing"] not in ("linspace", "leading", "trailing") or c["final_sigmas_type"] != "zero":
            raise ValueError("DPM++ 2M: timestep_spacing must be linspace / leading / trailing, final sigma zero")

    def _spaced_timesteps(self, n: int) -> np.ndarray:
        """diffusers 0.27.2 DPMSolverMultistepScheduler / UniPCMultistepScheduler.set_timesteps (lambda_min_clipped = -inf,
        so last_timestep = num_train_timesteps).  The SDXL scheduler config sets timestep_spacing="leading" explicitly
        (scripts/convert_from_A1111.py:947-959), and `from_config` carries it over to these classes."""
        c = self.config
        ntrain = c["num_train_timesteps"]
        sp = c["timestep_spacing"]
        if sp == "linspace":
            return np.linspace(0, ntrain - 1, n + 1).round()[::-1][:-1].copy().astype(np.int64)
        if sp == "leading":
            step_ratio = ntrain // (n + 1)
            ts = (np.arange(0, n + 1) * step_ratio).round()[::-1][:-1].copy().astype(np.int64)
            return ts + c["steps_offset"]
        if sp == "trailing":
            step_ratio = ntrain / n
            return np.arange(ntrain, 0, -step_ratio).round().copy().astype(np.int64) - 1
        raise ValueError(f"{sp} is not supported")

    @staticmethod
    def _sigma_to_t(sigma, log_sigmas):
        log_sigma = np.log(np.maximum(sigma, 1e-10))
        dists = log_sigma - log_sigmas[:, np.newaxis]
        low_idx = np.cumsum((dists >= 0), axis=0).argmax(axis=0).clip(max=log_sigmas.shape[0] - 2)
        high_idx = low_idx + 1
        low, high = log_sigmas[low_idx], log_sigmas[high_idx]
        w = np.clip((low - log_sigma) / (low - high), 0, 1)
        t = (1 - w) * low_idx + w * high_idx
        return t.reshape(np.shape(sigma))

    def set_timesteps(self, num_inference_steps, device=None):
        c = self.config
        ntrain = c["num_train_timesteps"]
        n = self.num_inference_steps = int(num_inference_steps)
        ts = self._spaced_timesteps(n)
        sig = (((1 - self.alphas_cumprod) / self.alphas_cumprod) ** 0.5).numpy()
        log_sig = np.log(sig)
        if c["use_karras_sigmas"]:
            sflip = np.flip(sig).copy()
            smin, smax = sflip[-1].item(), sflip[0].item()
            rho = 7.0
            ramp = np.linspace(0, 1, n)
            sig = (smax ** (1 / rho) + ramp * (smin ** (1 / rho) - smax ** (1 / rho))) ** rho
            ts = np.array([self._sigma_to_t(s, log_sig) for s in sig]).round()
        else:
            sig = np.interp(ts, np.arange(0, len(sig)), sig)
        sig = np.concatenate([sig, [0.0]]).astype(np.float32)
        self.sigmas = torch.from_numpy(sig)
        self._timesteps_host = ts.astype(np.int64)
        self._finish_set_timesteps(device)
        self.timesteps = torch.from_numpy(self._timesteps_host).to(self.device)

    @staticmethod
    def _alpha_sigma(sigma):
        alpha = 1 / ((sigma ** 2 + 1) ** 0.5)
        return alpha, sigma * alpha

    def _build_table(self):
        n = self.num_inference_steps
        tab = torch.zeros(n, N.STEP_ROW, dtype=torch.float32)
        for i in range(n):
            a_s, s_s = self._alpha_sigma(self.sigmas[i])
            a_t, s_t = self._alpha_sigma(self.sigmas[i + 1])
            lam_t = torch.log(a_t) - torch.log(s_t)
            lam_s = torch.log(a_s) - torch.log(s_s)
            h = lam_t - lam_s
            coef = a_t * (torch.exp(-h) - 1.0)
            tab[i, N.ST_TIMESTEP] = float(self._timesteps_host[i])
            tab[i, N.ST_IN_SCALE] = 1.0
            tab[i, N.ST_K1], tab[i, N.ST_K2] = s_s, 1.0 / a_s       # x0 = (x - sigma_hat*eps)/alpha
            tab[i, N.ST_CX] = s_t / s_s
            tab[i, N.ST_C0] = -coef
            first_order = (i == 0) or (i == n - 1)                  # lower_order_nums < 1, or final step (sigma -> 0)
            if not first_order:
                a_s1, s_s1 = self._alpha_sigma(self.sigmas[i - 1])
                lam_s1 = torch.log(a_s1) - torch.log(s_s1)
                r0 = (lam_s - lam_s1) / h
                tab[i, N.ST_C1] = -0.5 * coef * (1.0 / r0)
        return tab

    def noise_coeffs(self, timestep):
        a, s = self._alpha_sigma(self.sigmas[self.index_for_timestep(timestep)])
        return float(a), float(s)


class B200EulerAncestralDiscreteScheduler(B200EulerDiscreteScheduler):
    """EulerAncestralDiscreteScheduler (name `euler_a`, models/stable_diffusion.py:209-210): Euler to sigma_down,
    then fresh noise of scale sigma_up.  Same timestep / sigma tables as `euler`."""
    _stochastic = True

    def _build_table(self):
        tab = super()._build_table()
        for i in range(self.num_inference_steps):
            s_from, s_to = self.sigmas[i], self.sigmas[i + 1]
            s_up = (s_to ** 2 * (s_from ** 2 - s_to ** 2) / s_from ** 2) ** 0.5
            s_down = (s_to ** 2 - s_up ** 2) ** 0.5
            tab[i, N.ST_CE] = s_down - s_from
            tab[i, N.ST_CN] = s_up
        return tab


class B200DPMSolverSDEScheduler(B200DPMSolverMultistepScheduler):
    """`DPM++ 2M SDE Karras` as the reference builds it (models/stable_diffusion.py:215-218):
    algorithm_type="sde-dpmsolver++"; the Karras switch is spelt `se_karras_sigmas` there, which `from_config` drops,
    so the sigma spacing is the plain one — reproduced as is."""
    _stochastic = True
    _class_defaults = dict(B200DPMSolverMultistepScheduler._class_defaults, algorithm_type="sde-dpmsolver++")

    def __init__(self, se_karras_sigmas=None, **config):
        config.setdefault("algorithm_type", "sde-dpmsolver++")
        _SchedulerBase.__init__(self, **config)
        c = self.config
        if c["algorithm_type"] != "sde-dpmsolver++" or c["solver_order"] != 2 or c["solver_type"] != "midpoint":
            raise ValueError("only sde-dpmsolver++ / order 2 / midpoint is supported")
        if c["timestep_spacing"] not in ("linspace", "leading", "trailing") or c["final_sigmas_type"] != "zero":
            raise ValueError("DPM++ 2M SDE: timestep_spacing must be linspace / leading / trailing, final sigma zero")

    def _build_table(self):
        n = self.num_inference_steps
        tab = torch.zeros(n, N.STEP_ROW, dtype=torch.float32)
        for i in range(n):
            a_s, s_s = self._alpha_sigma(self.sigmas[i])
            a_t, s_t = self._alpha_sigma(self.sigmas[i + 1])
            lam_t = torch.log(a_t) - torch.log(s_t)
            lam_s = torch.log(a_s) - torch.log(s_s)
            h = lam_t - lam_s
            gain = a_t * (1 - torch.exp(-2.0 * h))
            tab[i, N.ST_TIMESTEP] = float(self._timesteps_host[i])
            tab[i, N.ST_IN_SCALE] = 1.0
            tab[i, N.ST_K1], tab[i, N.ST_K2] = s_s, 1.0 / a_s
            tab[i, N.ST_CX] = s_t / s_s * torch.exp(-h)
            tab[i, N.ST_C0] = gain
            tab[i, N.ST_CN] = s_t * torch.sqrt(1.0 - torch.exp(-2.0 * h))
            if not ((i == 0) or (i == n - 1)):
                a_s1, s_s1 = self._alpha_sigma(self.sigmas[i - 1])
                lam_s1 = torch.log(a_s1) - torch.log(s_s1)
                r0 = (lam_s - lam_s1) / h
                tab[i, N.ST_C1] = 0.5 * gain * (1.0 / r0)
        return tab


class B200PNDMScheduler(_SchedulerBase):
    """PNDMScheduler (name `PNDM`, models/stable_diffusion.py:219-220) with skip_prk_steps (all SD checkpoints): PLMS —
    linear multistep on eps (up to 4 terms), N + 1 UNet evaluations, the second timestep visited twice, the second
    evaluation redoing the first step from the remembered sample with the averaged eps."""
    _class_defaults = dict(timestep_spacing="leading", skip_prk_steps=True, set_alpha_to_one=False, steps_offset=0)
    init_noise_sigma = 1.0
    _uses_eps_history = True

    def __init__(self, **config):
        super().__init__(**config)
        if not self.config["skip_prk_steps"]:
            raise ValueError("PNDM with Runge-Kutta warm-up steps is not on the SD path (SD configs set skip_prk_steps)")
        self.final_alpha_cumprod = torch.tensor(1.0) if self.config["set_alpha_to_one"] else self.alphas_cumprod[0]
        self._start = 0

    def set_timesteps(self, num_inference_steps, device=None):
        c = self.config
        ntrain = c["num_train_timesteps"]
        n = self.num_inference_steps = int(num_inference_steps)
        sp = c["timestep_spacing"]
        if sp == "linspace":
            ts = np.linspace(0, ntrain - 1, n).round().astype(np.int64)
        elif sp == "leading":
            ts = (np.arange(0, n) * (ntrain // n)).round()
            ts += c["steps_offset"]
        elif sp == "trailing":
            ts = np.round(np.arange(ntrain, 0, -ntrain / n))[::-1].astype(np.int64)
            ts -= 1
        else:
            raise ValueError(f"{sp} is not supported")
        plms = np.concatenate([ts[:-1], ts[-2:-1], ts[-1:]])[::-1].copy()
        self._timesteps_host = plms.astype(np.int64)
        self._start = 0
        self._finish_set_timesteps(device)
        self.timesteps = torch.from_numpy(self._timesteps_host).to(self.device)

    def index_for_timestep(self, timestep) -> int:
        idx = np.nonzero(self._timesteps_host == int(timestep))[0]
        if len(idx) == 0:
            raise ValueError(f"timestep {timestep} is not in the schedule")
        return int(idx[0])

    def _build_table(self):
        return self._build_table_from(0)

    def _build_table_from(self, start: int):
        """Rows for the sequence executed from schedule position `start`: PNDM's special cases follow ITS call counter
        (0 on the first call), not the position in the schedule — exactly what diffusers does when img2img starts late."""
        ts = self._timesteps_host[start:]
        ratio = self.config["num_train_timesteps"] // self.num_inference_steps
        tab = torch.zeros(len(ts), N.STEP_ROW, dtype=torch.float32)
        n_ets = 0
        for counter, t in enumerate(int(v) for v in ts):
            prev_t = t - ratio
            if counter != 1:
                n_ets = min(n_ets, 3) + 1
            else:
                prev_t, t_used = t, t + ratio
            t_eff = t if counter != 1 else t_used
            a_t = self.alphas_cumprod[t_eff]
            a_p = self.alphas_cumprod[prev_t] if prev_t >= 0 else self.final_alpha_cumprod
            b_t, b_p = 1 - a_t, 1 - a_p
            sample_coeff = (a_p / a_t) ** 0.5
            g = -(a_p - a_t) / (a_t * b_p ** 0.5 + (a_t * b_t * a_p) ** 0.5)
            if n_ets == 1 and counter == 0:
                w, flags = (1.0, 0.0, 0.0, 0.0), N.STF_PUSH_EPS | N.STF_SAVE_X
            elif n_ets == 1 and counter == 1:
                w, flags = (0.5, 0.5, 0.0, 0.0), N.STF_USE_SAVED_X
            elif n_ets == 2:
                w, flags = (1.5, -0.5, 0.0, 0.0), N.STF_PUSH_EPS
            elif n_ets == 3:
                w, flags = (23 / 12, -16 / 12, 5 / 12, 0.0), N.STF_PUSH_EPS
            else:
                w, flags = (55 / 24, -59 / 24, 37 / 24, -9 / 24), N.STF_PUSH_EPS
            tab[counter, N.ST_TIMESTEP] = float(t)
            tab[counter, N.ST_IN_SCALE] = 1.0
            tab[counter, N.ST_K2] = 1.0
            tab[counter, N.ST_CX] = sample_coeff
            tab[counter, N.ST_CE], tab[counter, N.ST_H1] = g * w[0], g * w[1]
            tab[counter, N.ST_H2], tab[counter, N.ST_H3] = g * w[2], g * w[3]
            tab[counter, N.ST_FLAGS] = float(flags)
        return tab

    def rows_for(self, start_index: int, n: int) -> torch.Tensor:
        return self._build_table_from(start_index)[:n]

    # un-fused calls: the table is rebuilt from wherever the first step() lands
    def _init_step_index(self, timestep):
        if self._step_index is None:
            self._step_index = self._start = self.index_for_timestep(timestep)
            self._programs.clear()

    def _unfused_table(self, dev):
        key = f"pndm{self._start}@{dev}"
        if key not in self._dev:
            full = torch.zeros(len(self._table), N.STEP_ROW, dtype=torch.float32)
            rows = self._build_table_from(self._start)
            full[:len(rows)] = rows
            self._dev[key] = full.to(dev)
        return self._dev[key]

    def _unfused_counter(self) -> int:
        return self._step_index - self._start

    def noise_coeffs(self, timestep):
        a = self.alphas_cumprod[int(timestep)]
        return float(a ** 0.5), float((1 - a) ** 0.5)


class B200UniPCMultistepScheduler(B200DPMSolverMultistepScheduler):
    """UniPCMultistepScheduler (name `uni_pc`, models/stable_diffusion.py:221-222): predictor-corrector of order 2
    (bh2, predict_x0, lower_order_final), same timestep / sigma tables as DPM++ 2M.  Each call first CORRECTS the sample
    it is handed (UniC, using the fresh model output), then predicts the next one (UniP) — both are linear in
    {last corrected sample, the x0 predictions of this and the two previous calls}, so a row of coefficients drives the
    step kernel (B200SD_STF_UNIPC)."""
    _class_defaults = dict(timestep_spacing="linspace", solver_order=2, solver_type="bh2", predict_x0=True,
                           lower_order_final=True, use_karras_sigmas=False, final_sigmas_type="zero", steps_offset=0,
                           disable_corrector=[])
    _uses_history = True
    _uses_eps_history = True

    def __init__(self, **config):
        _SchedulerBase.__init__(self, **config)
        c = self.config
        if c["solver_order"] != 2 or c["solver_type"] != "bh2" or not c["predict_x0"] or not c["lower_order_final"]:
            raise ValueError("UniPC: only order 2 / bh2 / predict_x0 / lower_order_final (the class defaults) are supported")
        if (c["timestep_spacing"] not in ("linspace", "leading", "trailing") or c["final_sigmas_type"] != "zero"
                or c["disable_corrector"]):
            raise ValueError("UniPC: timestep_spacing must be linspace / leading / trailing, final sigma zero, corrector on")

    def _lam(self, sigma):
        a, s = self._alpha_sigma(sigma)
        return a, s, torch.log(a) - torch.log(s)

    def _build_table(self):
        return self._build_table_from(0)

    def _build_table_from(self, start: int):
        n_total = self.num_inference_steps
        n = n_total - start
        tab = torch.zeros(n, N.STEP_ROW, dtype=torch.float32)
        prev_order = 0
        for j in range(n):
            i = start + j                                   # position in the schedule; j = call counter
            a_s0, s_s0, lam_s0 = self._lam(self.sigmas[i])
            a_t, s_t, lam_t = self._lam(self.sigmas[i + 1])
            flags = N.STF_UNIPC
            tab[j, N.ST_TIMESTEP] = float(self._timesteps_host[i])
            tab[j, N.ST_IN_SCALE] = 1.0
            tab[j, N.ST_K1], tab[j, N.ST_K2] = s_s0, 1.0 / a_s0
            # ---- UniC: corrects the sample predicted by the previous call, order = that call's order ----
            if j > 0:
                q = prev_order
                a_c, s_c, lam_c = a_s0, s_s0, lam_s0            # "t" of the corrector is the current position
                a_p, s_p, lam_p = self._lam(self.sigmas[i - 1])  # its "s0" the previous one
                h = lam_c - lam_p
                hh = -h
                h_phi_1 = torch.expm1(hh)
                B_h = torch.expm1(hh)
                h_phi_k = h_phi_1 / hh - 1
                fact = 1
                b = []
                for k in range(1, q + 1):
                    b.append(h_phi_k * fact / B_h)
                    fact *= k + 1
                    h_phi_k = h_phi_k / hh - 1 / fact
                tab[j, N.ST_H1] = s_c / s_p
                tab[j, N.ST_H2] = -a_c * h_phi_1
                if q == 1:
                    rho_t = torch.tensor(0.5)
                else:
                    _, _, lam_p2 = self._lam(self.sigmas[i - 2])
                    rk = (lam_p2 - lam_p) / h
                    R = torch.stack([torch.ones(2), torch.stack([rk, torch.tensor(1.0)])])
                    rhos = torch.linalg.solve(R, torch.stack(b))
                    tab[j, N.ST_H3] = -a_c * B_h * rhos[0] / rk
                    rho_t = rhos[1]
                tab[j, N.ST_CE] = -a_c * B_h * rho_t
                flags |= N.STF_UNIPC_CORRECT
            # ---- UniP: predicts the next sample ----
            order = min(min(2, n_total - i), min(j, 2) + 1)
            h = lam_t - lam_s0
            hh = -h
            h_phi_1 = torch.expm1(hh)
            B_h = torch.expm1(hh)
            tab[j, N.ST_CX] = s_t / s_s0
            tab[j, N.ST_C0] = -a_t * h_phi_1
            if order == 2:
                _, _, lam_s1 = self._lam(self.sigmas[i - 1])
                rk = (lam_s1 - lam_s0) / h
                tab[j, N.ST_C1] = a_t * B_h * 0.5 / rk
            tab[j, N.ST_FLAGS] = float(flags)
            prev_order = order
        return tab

    def rows_for(self, start_index: int, n: int) -> torch.Tensor:
        return self._build_table_from(start_index)[:n]

    # un-fused calls: like PNDM, the warm-up follows the call counter
    def _init_step_index(self, timestep):
        if self._step_index is None:
            self._step_index = self._start = self.index_for_timestep(timestep)
            self._programs.clear()

    def _unfused_table(self, dev):
        key = f"unipc{self._start}@{dev}"
        if key not in self._dev:
            full = torch.zeros(len(self._table), N.STEP_ROW, dtype=torch.float32)
            rows = self._build_table_from(self._start)
            full[:len(rows)] = rows
            self._dev[key] = full.to(dev)
        return self._dev[key]

    def _unfused_counter(self) -> int:
        return self._step_index - self._start


SCHEDULER_CLASSES = {
    "euler": (B200EulerDiscreteScheduler, {}),
    "DDIM": (B200DDIMScheduler, {}),
    "DPM++ 2M": (B200DPMSolverMultistepScheduler, {}),
    "DPM++ 2M Karras": (B200DPMSolverMultistepScheduler, dict(use_karras_sigmas=True)),
    "euler_a": (B200EulerAncestralDiscreteScheduler, {}),
    # the reference's own spelling of the call, typo included (models/stable_diffusion.py:215-218)
    "DPM++ 2M SDE Karras": (B200DPMSolverSDEScheduler, dict(se_karras_sigmas=True, algorithm_type="sde-dpmsolver++")),
    "PNDM": (B200PNDMScheduler, {}),
    "uni_pc": (B200UniPCMultistepScheduler, {}),
}
# every name the reference accepts (models/stable_diffusion.py:205-222) has a kernel path now
NEXT_ROW_SCHEDULERS = ()


def make_scheduler(name: str, config) -> _SchedulerBase:
    if name in SCHEDULER_CLASSES:
        cls, extra = SCHEDULER_CLASSES[name]
        return cls.from_config(config, **extra)
    if name in NEXT_ROW_SCHEDULERS:
        raise ValueError(f"scheduler '{name}' has no B200 kernel yet (no CPU fallback is provided)")
    raise ValueError(f"Unknown scheduler name: {name}")
