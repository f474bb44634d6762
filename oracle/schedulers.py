"""Restatement of the diffusers 0.27.2 schedulers the reference can select on this path
(models/stable_diffusion.py:135-138 default Euler; :205-214 DDIM / euler / DPM++ 2M / DPM++ 2M Karras).

Test infrastructure (see oracle/__init__.py).  Table construction repeats diffusers' numpy expressions so
`timesteps` / `sigmas` are reproduced bit for bit; `step` keeps diffusers' operation order
in fp32.  Scheduler hyper-parameters of SD checkpoints: scripts/convert_from_A1111.py:947-982.
"""
from __future__ import annotations

import numpy as np
import torch

SD15_SCHEDULER_CONFIG = dict(num_train_timesteps=1000, beta_start=0.00085, beta_end=0.012, beta_schedule="scaled_linear",
                             steps_offset=1, clip_sample=False, set_alpha_to_one=False, skip_prk_steps=True,
                             prediction_type="epsilon")
SDXL_SCHEDULER_CONFIG = dict(num_train_timesteps=1000, beta_start=0.00085, beta_end=0.012, beta_schedule="scaled_linear",
                             steps_offset=1, timestep_spacing="leading", interpolation_type="linear",
                             use_karras_sigmas=False, prediction_type="epsilon", clip_sample=False,
                             set_alpha_to_one=False, skip_prk_steps=True, sample_max_value=1.0)


def _betas(cfg):
    n = cfg["num_train_timesteps"]
    if cfg.get("beta_schedule", "linear") == "scaled_linear":
        return torch.linspace(cfg["beta_start"] ** 0.5, cfg["beta_end"] ** 0.5, n, dtype=torch.float32) ** 2
    return torch.linspace(cfg["beta_start"], cfg["beta_end"], n, dtype=torch.float32)


class _Base:
    order = 1

    def __init__(self, **config):
        self.config = dict(config)
        self.betas = _betas(self.config)
        self.alphas = 1.0 - self.betas
        self.alphas_cumprod = torch.cumprod(self.alphas, dim=0)
        self.timesteps = None
        self.sigmas = None
        self.step_index = None
        self.num_inference_steps = None

    def _index_for_timestep(self, t):
        idx = (self.timesteps == torch.as_tensor(t).to(self.timesteps.dtype)).nonzero()
        pos = 1 if len(idx) > 1 else 0
        return idx[pos].item()

    def _init_step_index(self, t):
        if self.step_index is None:
            self.step_index = self._index_for_timestep(t)


class EulerDiscreteOracle(_Base):
    def __init__(self, **config):
        super().__init__(**config)
        self.spacing = self.config.get("timestep_spacing", "linspace")
        sig = ((1 - self.alphas_cumprod) / self.alphas_cumprod) ** 0.5
        self.sigmas = torch.cat([sig.flip(0), torch.zeros(1)])

    @property
    def init_noise_sigma(self):
        m = self.sigmas.max()
        if self.spacing in ("linspace", "trailing"):
            return m
        return (m ** 2 + 1) ** 0.5

    def set_timesteps(self, n, device=None):
        N = self.config["num_train_timesteps"]
        self.num_inference_steps = n
        if self.spacing == "linspace":
            ts = np.linspace(0, N - 1, n, dtype=np.float32)[::-1].copy()
        elif self.spacing == "leading":
            ts = (np.arange(0, n) * (N // n)).round()[::-1].copy().astype(np.float32)
            ts += self.config.get("steps_offset", 0)
        else:
            ts = (np.arange(N, 0, -N / n)).round().copy().astype(np.float32)
            ts -= 1
        sig = (((1 - self.alphas_cumprod) / self.alphas_cumprod) ** 0.5).numpy()
        sig = np.interp(ts, np.arange(0, len(sig)), sig)
        sig = torch.from_numpy(sig).to(dtype=torch.float32)
        self.timesteps = torch.from_numpy(ts.astype(np.float32))
        self.sigmas = torch.cat([sig, torch.zeros(1)])
        self.step_index = None

    def scale_model_input(self, x, t):
        self._init_step_index(t)
        s = self.sigmas[self.step_index]
        return x / ((s ** 2 + 1) ** 0.5)

    def step(self, eps, t, x):
        self._init_step_index(t)
        x = x.to(torch.float32)
        s = self.sigmas[self.step_index]
        x0 = x - s * eps
        d = (x - x0) / s
        dt = self.sigmas[self.step_index + 1] - s
        self.step_index += 1
        return x + d * dt

    def add_noise(self, x0, noise, timesteps):
        idx = [self._index_for_timestep(t) for t in timesteps]
        s = self.sigmas[idx].flatten().to(x0.dtype)
        while s.dim() < x0.dim():
            s = s.unsqueeze(-1)
        return x0 + noise * s


class DDIMOracle(_Base):
    init_noise_sigma = 1.0

    def __init__(self, **config):
        super().__init__(**config)
        self.final_alpha_cumprod = torch.tensor(1.0) if self.config.get("set_alpha_to_one", True) else self.alphas_cumprod[0]

    def set_timesteps(self, n, device=None):
        N = self.config["num_train_timesteps"]
        self.num_inference_steps = n
        spacing = self.config.get("timestep_spacing", "leading")
        if spacing == "leading":
            ts = (np.arange(0, n) * (N // n)).round()[::-1].copy().astype(np.int64)
            ts += self.config.get("steps_offset", 0)
        elif spacing == "linspace":
            ts = np.linspace(0, N - 1, n).round()[::-1].copy().astype(np.int64)
        else:
            ts = np.round(np.arange(N, 0, -N / n)).astype(np.int64)
            ts -= 1
        self.timesteps = torch.from_numpy(ts)

    def scale_model_input(self, x, t):
        return x

    def step(self, eps, t, x):
        t = int(t)
        prev_t = t - self.config["num_train_timesteps"] // self.num_inference_steps
        a_t = self.alphas_cumprod[t]
        a_p = self.alphas_cumprod[prev_t] if prev_t >= 0 else self.final_alpha_cumprod
        b_t = 1 - a_t
        x0 = (x - b_t ** 0.5 * eps) / a_t ** 0.5
        if self.config.get("clip_sample", True):
            x0 = x0.clamp(-self.config.get("clip_sample_range", 1.0), self.config.get("clip_sample_range", 1.0))
        direction = (1 - a_p) ** 0.5 * eps
        return a_p ** 0.5 * x0 + direction

    def add_noise(self, x0, noise, timesteps):
        a = self.alphas_cumprod[timesteps.long()].to(x0.dtype)
        sa, sb = a ** 0.5, (1 - a) ** 0.5
        while sa.dim() < x0.dim():
            sa, sb = sa.unsqueeze(-1), sb.unsqueeze(-1)
        return sa * x0 + sb * noise


class DPMSolverMultistepOracle(_Base):
    """DPM-Solver++(2M), midpoint, epsilon prediction, final sigma 0, optional Karras sigmas."""
    init_noise_sigma = 1.0

    def __init__(self, **config):
        super().__init__(**config)
        self.use_karras = bool(self.config.get("use_karras_sigmas", False))
        self.m = [None, None]
        self.lower_order_nums = 0

    def set_timesteps(self, n, device=None):
        N = self.config["num_train_timesteps"]
        self.num_inference_steps = n
        # diffusers 0.27.2 DPMSolverMultistepScheduler.set_timesteps (UniPC: the same three forms); lambda_min_clipped = -inf
        spacing = self.config.get("timestep_spacing", "linspace")
        if spacing == "linspace":
            ts = np.linspace(0, N - 1, n + 1).round()[::-1][:-1].copy().astype(np.int64)
        elif spacing == "leading":
            step_ratio = N // (n + 1)
            ts = (np.arange(0, n + 1) * step_ratio).round()[::-1][:-1].copy().astype(np.int64)
            ts += self.config.get("steps_offset", 0)
        elif spacing == "trailing":
            ts = np.arange(N, 0, -(N / n)).round().copy().astype(np.int64)
            ts -= 1
        else:
            raise ValueError(f"{spacing} is not supported")
        sig = (((1 - self.alphas_cumprod) / self.alphas_cumprod) ** 0.5).numpy()
        log_sig = np.log(sig)
        if self.use_karras:
            sflip = np.flip(sig).copy()
            smin, smax = sflip[-1].item(), sflip[0].item()
            ramp = np.linspace(0, 1, n)
            rho = 7.0
            sig = (smax ** (1 / rho) + ramp * (smin ** (1 / rho) - smax ** (1 / rho))) ** rho
            ts = np.array([self._sigma_to_t(s, log_sig) for s in sig]).round()
        else:
            sig = np.interp(ts, np.arange(0, len(sig)), sig)
        # final_sigmas_type: "zero" (class default, what every SD config gets) or "sigma_min" (only used to check this
        # restatement against diffusers' own scheduler unit tests, which pin the older behaviour)
        if self.config.get("final_sigmas_type", "zero") == "sigma_min":
            sigma_last = float(((1 - self.alphas_cumprod[0]) / self.alphas_cumprod[0]) ** 0.5)
        else:
            sigma_last = 0.0
        sig = np.concatenate([sig, [sigma_last]]).astype(np.float32)
        self.sigmas = torch.from_numpy(sig)
        self.timesteps = torch.from_numpy(ts).to(torch.int64)
        self.m = [None, None]
        self.lower_order_nums = 0
        self.step_index = None

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

    @staticmethod
    def _alpha_sigma(sigma):
        alpha = 1 / ((sigma ** 2 + 1) ** 0.5)
        return alpha, sigma * alpha

    def scale_model_input(self, x, t):
        return x

    def _lower_final(self, i, n):
        """diffusers: (step_index == len(timesteps) - 1) and (euler_at_final or (lower_order_final and len < 15) or
        final_sigmas_type == "zero") — with the class defaults the last step is always first-order."""
        c = self.config
        return i == n - 1 and (c.get("euler_at_final", False) or (c.get("lower_order_final", True) and n < 15)
                               or c.get("final_sigmas_type", "zero") == "zero")

    def step(self, eps, t, x):
        self._init_step_index(t)
        i = self.step_index
        n = len(self.timesteps)
        lower_final = self._lower_final(i, n)
        a_s, s_s = self._alpha_sigma(self.sigmas[i])
        x0 = (x - s_s * eps) / a_s
        self.m = [self.m[1], x0]
        x = x.to(torch.float32)
        a_t, s_t = self._alpha_sigma(self.sigmas[i + 1])
        lam_t = torch.log(a_t) - torch.log(s_t)
        lam_s = torch.log(a_s) - torch.log(s_s)
        h = lam_t - lam_s
        if self.lower_order_nums < 1 or lower_final:
            out = (s_t / s_s) * x - (a_t * (torch.exp(-h) - 1.0)) * x0
        else:
            a_s1, s_s1 = self._alpha_sigma(self.sigmas[i - 1])
            lam_s1 = torch.log(a_s1) - torch.log(s_s1)
            h0 = lam_s - lam_s1
            r0 = h0 / h
            D0, D1 = self.m[1], (1.0 / r0) * (self.m[1] - self.m[0])
            out = (s_t / s_s) * x - (a_t * (torch.exp(-h) - 1.0)) * D0 - 0.5 * (a_t * (torch.exp(-h) - 1.0)) * D1
        if self.lower_order_nums < 2:
            self.lower_order_nums += 1
        self.step_index += 1
        return out

    def add_noise(self, x0, noise, timesteps):
        idx = [self._index_for_timestep(t) for t in timesteps]
        a, s = self._alpha_sigma(self.sigmas[idx].flatten().to(x0.dtype))
        while a.dim() < x0.dim():
            a, s = a.unsqueeze(-1), s.unsqueeze(-1)
        return a * x0 + s * noise


class EulerAncestralOracle(EulerDiscreteOracle):
    """EulerAncestralDiscreteScheduler (`euler_a`, models/stable_diffusion.py:209-210).  The reference calls
    `scheduler.step(noise_pred, t, latents)` without a generator (pipeline :489), i.e. the noise comes from the global
    device RNG; the oracle takes the draw as an argument so that both sides can be fed the same one."""

    def step(self, eps, t, x, noise):
        self._init_step_index(t)
        x = x.to(torch.float32)
        sigma = self.sigmas[self.step_index]
        x0 = x - sigma * eps
        sigma_from, sigma_to = self.sigmas[self.step_index], self.sigmas[self.step_index + 1]
        sigma_up = (sigma_to ** 2 * (sigma_from ** 2 - sigma_to ** 2) / sigma_from ** 2) ** 0.5
        sigma_down = (sigma_to ** 2 - sigma_up ** 2) ** 0.5
        d = (x - x0) / sigma
        dt = sigma_down - sigma
        self.step_index += 1
        return x + d * dt + noise * sigma_up


class DPMSolverSDEOracle(DPMSolverMultistepOracle):
    """DPMSolverMultistepScheduler with algorithm_type="sde-dpmsolver++" (`DPM++ 2M SDE Karras`,
    models/stable_diffusion.py:215-218 — the reference spells the Karras switch `se_karras_sigmas`, which
    `from_config` drops, so the sigmas are NOT Karras-spaced; reproduced as is)."""

    def step(self, eps, t, x, noise):
        self._init_step_index(t)
        i = self.step_index
        n = len(self.timesteps)
        lower_final = self._lower_final(i, n)
        a_s, s_s = self._alpha_sigma(self.sigmas[i])
        x0 = (x - s_s * eps) / a_s
        self.m = [self.m[1], x0]
        x = x.to(torch.float32)
        a_t, s_t = self._alpha_sigma(self.sigmas[i + 1])
        lam_t = torch.log(a_t) - torch.log(s_t)
        lam_s = torch.log(a_s) - torch.log(s_s)
        h = lam_t - lam_s
        if self.lower_order_nums < 1 or lower_final:
            out = (s_t / s_s * torch.exp(-h)) * x + (a_t * (1 - torch.exp(-2.0 * h))) * x0 \
                + s_t * torch.sqrt(1.0 - torch.exp(-2.0 * h)) * noise
        else:
            a_s1, s_s1 = self._alpha_sigma(self.sigmas[i - 1])
            lam_s1 = torch.log(a_s1) - torch.log(s_s1)
            r0 = (lam_s - lam_s1) / h
            D0, D1 = self.m[1], (1.0 / r0) * (self.m[1] - self.m[0])
            out = (s_t / s_s * torch.exp(-h)) * x + (a_t * (1 - torch.exp(-2.0 * h))) * D0 \
                + 0.5 * (a_t * (1 - torch.exp(-2.0 * h))) * D1 + s_t * torch.sqrt(1.0 - torch.exp(-2.0 * h)) * noise
        if self.lower_order_nums < 2:
            self.lower_order_nums += 1
        self.step_index += 1
        return out


class PNDMOracle(_Base):
    """PNDMScheduler (`PNDM`, models/stable_diffusion.py:219-220) with skip_prk_steps=True (every SD checkpoint:
    scripts/convert_from_A1111.py:947-982) — i.e. PLMS: N + 1 UNet evaluations, the second timestep visited twice."""
    init_noise_sigma = 1.0

    def __init__(self, **config):
        super().__init__(**config)
        if not self.config.get("skip_prk_steps", False):
            raise ValueError("PNDM with Runge-Kutta warm-up steps is not used by SD checkpoints")
        self.final_alpha_cumprod = torch.tensor(1.0) if self.config.get("set_alpha_to_one", False) else self.alphas_cumprod[0]
        self.spacing = self.config.get("timestep_spacing", "leading")

    def set_timesteps(self, n, device=None):
        N = self.config["num_train_timesteps"]
        self.num_inference_steps = n
        if self.spacing == "linspace":
            ts = np.linspace(0, N - 1, n).round().astype(np.int64)
        elif self.spacing == "leading":
            ts = (np.arange(0, n) * (N // n)).round()
            ts += self.config.get("steps_offset", 0)
        else:
            ts = np.round(np.arange(N, 0, -N / n))[::-1].astype(np.int64)
            ts -= 1
        plms = np.concatenate([ts[:-1], ts[-2:-1], ts[-1:]])[::-1].copy()
        self.timesteps = torch.from_numpy(plms.astype(np.int64))
        self.ets, self.counter, self.cur_sample = [], 0, None

    def scale_model_input(self, x, t):
        return x

    def _prev(self, x, t, prev_t, mo):
        a_t = self.alphas_cumprod[t]
        a_p = self.alphas_cumprod[prev_t] if prev_t >= 0 else self.final_alpha_cumprod
        b_t, b_p = 1 - a_t, 1 - a_p
        sample_coeff = (a_p / a_t) ** 0.5
        denom = a_t * b_p ** 0.5 + (a_t * b_t * a_p) ** 0.5
        return sample_coeff * x - (a_p - a_t) * mo / denom

    def step(self, eps, t, x):
        t = int(t)
        ratio = self.config["num_train_timesteps"] // self.num_inference_steps
        prev_t = t - ratio
        x = x.to(torch.float32)
        if self.counter != 1:
            self.ets = self.ets[-3:]
            self.ets.append(eps)
        else:
            prev_t = t
            t = t + ratio
        if len(self.ets) == 1 and self.counter == 0:
            mo = eps
            self.cur_sample = x
        elif len(self.ets) == 1 and self.counter == 1:
            mo = (eps + self.ets[-1]) / 2
            x = self.cur_sample
            self.cur_sample = None
        elif len(self.ets) == 2:
            mo = (3 * self.ets[-1] - self.ets[-2]) / 2
        elif len(self.ets) == 3:
            mo = (23 * self.ets[-1] - 16 * self.ets[-2] + 5 * self.ets[-3]) / 12
        else:
            mo = (1 / 24) * (55 * self.ets[-1] - 59 * self.ets[-2] + 37 * self.ets[-3] - 9 * self.ets[-4])
        out = self._prev(x, t, prev_t, mo)
        self.counter += 1
        return out

    def add_noise(self, x0, noise, timesteps):
        a = self.alphas_cumprod[timesteps.long()].flatten().to(x0.dtype)
        while a.dim() < x0.dim():
            a = a.unsqueeze(-1)
        return a ** 0.5 * x0 + (1 - a) ** 0.5 * noise


class UniPCOracle(DPMSolverMultistepOracle):
    """UniPCMultistepScheduler (`uni_pc`, models/stable_diffusion.py:221-222), class defaults: solver_order 2, bh2,
    predict_x0, lower_order_final, corrector on every step after the first, final sigma zero.  Restated in diffusers'
    structure: `step` = [UniC on the incoming sample] -> shift the x0 history -> UniP."""

    def set_timesteps(self, n, device=None):
        super().set_timesteps(n, device)
        self.model_outputs = [None, None]
        self.lower_order_nums = 0
        self.last_sample = None
        self.this_order = None

    def _lam(self, sigma):
        a, s = self._alpha_sigma(sigma)
        return a, s, torch.log(a) - torch.log(s)

    def _b_coeffs(self, hh, order):
        h_phi_1 = torch.expm1(hh)
        B_h = torch.expm1(hh)  # bh2
        h_phi_k = h_phi_1 / hh - 1
        factorial_i = 1
        b = []
        for i in range(1, order + 1):
            b.append(h_phi_k * factorial_i / B_h)
            factorial_i *= i + 1
            h_phi_k = h_phi_k / hh - 1 / factorial_i
        return h_phi_1, B_h, torch.stack(b)

    def _uni_p(self, sample, order):
        m0 = self.model_outputs[-1]
        x = sample
        a_t, s_t, lam_t = self._lam(self.sigmas[self.step_index + 1])
        a_s0, s_s0, lam_s0 = self._lam(self.sigmas[self.step_index])
        h = lam_t - lam_s0
        rks, D1s = [], []
        for i in range(1, order):
            si = self.step_index - i
            mi = self.model_outputs[-(i + 1)]
            _, _, lam_si = self._lam(self.sigmas[si])
            rk = (lam_si - lam_s0) / h
            rks.append(rk)
            D1s.append((mi - m0) / rk)
        h_phi_1, B_h, b = self._b_coeffs(-h, order)
        x_t_ = s_t / s_s0 * x - a_t * h_phi_1 * m0
        pred_res = 0.5 * D1s[0] if D1s else 0  # order 2: rhos_p = [0.5]
        return x_t_ - a_t * B_h * pred_res

    def _uni_c(self, this_model_output, last_sample, order):
        m0 = self.model_outputs[-1]
        x = last_sample
        model_t = this_model_output
        a_t, s_t, lam_t = self._lam(self.sigmas[self.step_index])
        a_s0, s_s0, lam_s0 = self._lam(self.sigmas[self.step_index - 1])
        h = lam_t - lam_s0
        rks, D1s = [], []
        for i in range(1, order):
            si = self.step_index - (i + 1)
            mi = self.model_outputs[-(i + 1)]
            _, _, lam_si = self._lam(self.sigmas[si])
            rk = (lam_si - lam_s0) / h
            rks.append(rk)
            D1s.append((mi - m0) / rk)
        rks.append(torch.tensor(1.0))
        rks = torch.stack(rks)
        h_phi_1, B_h, b = self._b_coeffs(-h, order)
        R = torch.stack([torch.pow(rks, i - 1) for i in range(1, order + 1)])
        rhos_c = torch.tensor([0.5]) if order == 1 else torch.linalg.solve(R, b)
        x_t_ = s_t / s_s0 * x - a_t * h_phi_1 * m0
        corr_res = sum(rhos_c[k] * D1s[k] for k in range(len(D1s))) if D1s else 0
        D1_t = model_t - m0
        return x_t_ - a_t * B_h * (corr_res + rhos_c[-1] * D1_t)

    def step(self, eps, t, x):
        self._init_step_index(t)
        i = self.step_index
        x = x.to(torch.float32)
        use_corrector = i > 0 and self.last_sample is not None
        a_s, s_s = self._alpha_sigma(self.sigmas[i])
        x0 = (x - s_s * eps) / a_s                                   # convert_model_output (epsilon, predict_x0)
        if use_corrector:
            x = self._uni_c(x0, self.last_sample, self.this_order)
        self.model_outputs = [self.model_outputs[1], x0]
        this_order = min(2, len(self.timesteps) - i) if self.config.get("lower_order_final", True) else 2
        self.this_order = min(this_order, self.lower_order_nums + 1)  # warm-up
        self.last_sample = x
        out = self._uni_p(x, self.this_order)
        if self.lower_order_nums < 2:
            self.lower_order_nums += 1
        self.step_index += 1
        return out


STOCHASTIC = ("euler_a", "DPM++ 2M SDE Karras")


def make_scheduler(name: str, config: dict):
    """Name table of models/stable_diffusion.py:205-222."""
    if name == "euler":
        return EulerDiscreteOracle(**config)
    if name == "DDIM":
        return DDIMOracle(**config)
    if name == "DPM++ 2M":
        return DPMSolverMultistepOracle(**config)
    if name == "DPM++ 2M Karras":
        return DPMSolverMultistepOracle(**dict(config, use_karras_sigmas=True))
    if name == "euler_a":
        return EulerAncestralOracle(**config)
    if name == "DPM++ 2M SDE Karras":  # (`se_karras_sigmas=True` of the reference is dropped by from_config: not Karras)
        return DPMSolverSDEOracle(**config)
    if name == "PNDM":
        return PNDMOracle(**config)
    if name == "uni_pc":
        return UniPCOracle(**config)
    raise ValueError(f"Unknown scheduler name: {name}")
