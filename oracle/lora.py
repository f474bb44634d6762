"""Unmerged (PEFT-semantics) LoRA forward on the oracle modules — test infrastructure.
y = W x + sum_i w_i * (alpha_i / r_i) * B_i (A_i x)   (models/stable_diffusion.py:298-309)."""
from __future__ import annotations

import torch
import torch.nn as nn


def attach_unmerged(model: nn.Module, adapters: dict, weights: dict):
    """adapters: {name: {module_path: {"down": A, "up": B, "alpha": float|None}}}.  Returns the hook handles."""
    handles = []
    mods = dict(model.named_modules())
    for name, layers in adapters.items():
        w = float(weights.get(name, 1.0))
        for path, ent in layers.items():
            m = mods[path]
            A, B = ent["down"].float(), ent["up"].float()
            scale = (ent["alpha"] / A.shape[0]) if ent.get("alpha") is not None else 1.0

            def hook(mod, inp, out, A=A, B=B, s=w * scale):
                x = inp[0]
                return out + s * (x @ A.to(x.device, x.dtype).t() @ B.to(x.device, x.dtype).t())

            handles.append(m.register_forward_hook(hook))
    return handles
