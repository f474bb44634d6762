"""Random-init weights of the exact named architectures (no network / checkpoints are available):
N(0, 1/fan_in) for conv / linear weights, small N(0, 0.02) biases, norm gamma ~ 1, beta ~ 0.
Deterministic for a given seed; keys and shapes are diffusers' so the same dict loads into the oracle."""
from __future__ import annotations

import math
from collections import OrderedDict
from typing import Dict

import torch


def random_state_dict(shapes: "OrderedDict[str, tuple]", seed: int = 0, dtype=torch.float32, gain: float = 1.0) -> Dict[str, torch.Tensor]:
    g = torch.Generator(device="cpu").manual_seed(seed)
    sd: Dict[str, torch.Tensor] = OrderedDict()
    for name, shp in shapes.items():
        leaf = name.rsplit(".", 2)
        is_norm = any(part.startswith("norm") or part in ("conv_norm_out", "group_norm") for part in name.split(".")[-2:-1])
        if name.endswith(".weight") and len(shp) >= 2:
            fan_in = 1
            for d in shp[1:]:
                fan_in *= d
            t = torch.randn(shp, generator=g) * (gain / math.sqrt(fan_in))
        elif name.endswith(".weight") and is_norm:
            t = 1.0 + 0.1 * torch.randn(shp, generator=g)
        elif name.endswith(".bias") and is_norm:
            t = 0.05 * torch.randn(shp, generator=g)
        elif name.endswith(".bias"):
            t = 0.02 * torch.randn(shp, generator=g)
        else:
            t = torch.randn(shp, generator=g)
        sd[name] = t.to(dtype)
        del leaf
    return sd
