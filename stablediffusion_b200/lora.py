"""LoRA handling for the hot path: parse the file formats the reference accepts and MERGE the adapters
into the base weights at load time (zero runtime cost).

Reference semantics being reproduced (models/stable_diffusion.py:230-309): several named adapters can be
loaded, `set_adapters(names, weights)` gives per-adapter weights w_i, and the effective layer is
    y = W x + sum_i w_i * (alpha_i / r_i) * B_i (A_i x)          (PEFT LoRA, diffusers 0.27.2)
which equals a layer with W' = W + sum_i w_i * (alpha_i / r_i) * B_i A_i.  The merge is done in fp32 on the
pristine weights, so set/delete/re-load is exact (no accumulated rounding).

Formats:
  * `train_lora_pipeline` output (pipelines/train_lora_pipeline.py:497-528 -> `save_lora_weights`):
    `unet.<module path>.{lora.down,lora.up}.weight` (also the PEFT spelling `lora_A/lora_B`), text encoders under
    `text_encoder[_2].…`; no alpha tensors (alpha = r -> scale 1).
  * A1111 / kohya: `lora_unet_<path with underscores>.{lora_down,lora_up}.weight` + `.alpha`.
"""
from __future__ import annotations

import re
from typing import Dict, Optional, Tuple

import torch

_DOWN = ("lora.down.weight", "lora_A.weight", "lora_down.weight", "lora_linear_layer.down.weight", "down.weight")
_UP = ("lora.up.weight", "lora_B.weight", "lora_up.weight", "lora_linear_layer.up.weight", "up.weight")


def _split_suffix(key: str) -> Optional[Tuple[str, str]]:
    for suf in sorted(_DOWN, key=len, reverse=True):
        if key.endswith("." + suf):
            return key[: -len(suf) - 1], "down"
    for suf in sorted(_UP, key=len, reverse=True):
        if key.endswith("." + suf):
            return key[: -len(suf) - 1], "up"
    if key.endswith(".alpha"):
        return key[:-6], "alpha"
    return None


def _kohya_unet_module(name: str, known_modules) -> Optional[str]:
    """lora_unet_down_blocks_0_attentions_0_transformer_blocks_0_attn1_to_q -> dotted diffusers module path."""
    body = name[len("lora_unet_"):]
    flat = {m.replace(".", "_"): m for m in known_modules}
    return flat.get(body)


def parse_lora_state_dict(sd: Dict[str, torch.Tensor], unet_modules=None) -> Dict[str, Dict[str, dict]]:
    """-> {"unet": {module_path: {"down": A [r,in], "up": B [out,r], "alpha": float|None}}, "text_encoder": {...}, ...}"""
    if not all("lora" in k for k in sd.keys()):
        raise ValueError("Invalid LoRA checkpoint.")  # models/stable_diffusion.py:264-266
    out: Dict[str, Dict[str, dict]] = {"unet": {}, "text_encoder": {}, "text_encoder_2": {}}
    unresolved = set()
    for key, val in sd.items():
        sp = _split_suffix(key)
        if sp is None:
            continue
        mod, kind = sp
        if mod.startswith("lora_unet_"):
            if unet_modules is None:
                raise ValueError("kohya-format LoRA needs the UNet module list to resolve names")
            path = _kohya_unet_module(mod, unet_modules)
            if path is None:  # e.g. SGM / LDM block names (lora_unet_input_blocks_*): not a diffusers module path
                unresolved.add(mod)
                continue
            part = "unet"
        elif mod.startswith("lora_te"):
            part = "text_encoder_2" if mod.startswith("lora_te2_") else "text_encoder"
            path = mod
        else:
            part = None
            for pfx in ("text_encoder_2.", "text_encoder.", "unet."):
                if mod.startswith(pfx):
                    part, path = pfx[:-1], mod[len(pfx):]
                    break
            if part is None:
                part, path = "unet", mod
            path = re.sub(r"\.processor", "", path)
            path = path.replace(".to_out_lora", ".to_out.0").replace(".to_q_lora", ".to_q")
            path = path.replace(".to_k_lora", ".to_k").replace(".to_v_lora", ".to_v")
        ent = out[part].setdefault(path, {"down": None, "up": None, "alpha": None})
        ent[kind] = float(val) if kind == "alpha" else val
    if unresolved:
        # a file whose modules cannot be placed must not load as a silent no-op adapter
        ex = sorted(unresolved)[:3]
        raise ValueError(f"LoRA checkpoint targets {len(unresolved)} UNet modules this model does not have, e.g. {ex} "
                         "(SGM/LDM-style block names are not supported: convert the file to diffusers/kohya-diffusers names)")
    for part in out.values():
        for path, ent in list(part.items()):
            if ent["down"] is None or ent["up"] is None:
                del part[path]
    return out


def lora_delta(ent: dict) -> torch.Tensor:
    """(alpha / r) * B A as an fp32 matrix shaped like the target weight (Linear, 1x1 or 3x3 conv)."""
    A, B = ent["down"].float(), ent["up"].float()
    r = A.shape[0]
    scale = (ent["alpha"] / r) if ent.get("alpha") is not None else 1.0
    if A.dim() == 4:  # conv LoRA: down [r, in, kh, kw], up [out, r, 1, 1]
        d = torch.einsum("or,rikl->oikl", B.reshape(B.shape[0], r), A)
    else:
        d = B @ A
    return d * scale


def merged_state_dict(master: Dict[str, torch.Tensor], adapters: Dict[str, dict], weights: Dict[str, float],
                      prefix: str = "unet") -> Dict[str, torch.Tensor]:
    """W' = W + sum_i w_i * delta_i for every targeted module; untouched tensors are passed through."""
    if not adapters:
        return master
    out = dict(master)
    for name, layers in adapters.items():
        w = float(weights.get(name, 1.0))
        if w == 0.0:
            continue
        for path, ent in layers.items():
            key = path + ".weight"
            if key not in master:
                raise ValueError(f"LoRA adapter '{name}' targets unknown module '{path}'")
            base = out[key].float()
            out[key] = base + w * lora_delta(ent).reshape(base.shape).to(base.device)
    return out
