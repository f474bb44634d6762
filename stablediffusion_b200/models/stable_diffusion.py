"""`SDModelWrapper` with the public surface of the reference's models/stable_diffusion.py:40-335, but whose
`.base`, `.vae` and `.scheduler` are the B200-native objects (B200UNet / B200VAE / B200*Scheduler).

Loading (reference :106-156 `load_hf_checkpoint`): an HF-layout folder — `unet/`, `vae/`, `text_encoder/`,
`tokenizer/`, `scheduler/` (+ `text_encoder_2/`, `tokenizer_2/` for SDXL) — with fp16-variant safetensors
(`diffusion_pytorch_model.fp16.safetensors`), i.e. exactly what scripts/convert_from_A1111.py:1381-1386 writes.
Weights are converted to bf16 and re-laid-out for the kernels at load.  There is no network in this build:
hub ids (`OnMoon/...`, `runwayml/...`) must already be local folders, otherwise loading raises.
Tokenizers stay transformers objects; the text encoders are read with transformers and become native
`B200CLIPTextModel`s when the wrapper moves to the device.
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Union

import torch

from ..config import UNET_DEFAULTS, VAE_DEFAULTS, Config, load_json_config, scheduler_config, unet_config, vae_config
from ..lora import parse_lora_state_dict
from ..clip import B200CLIPTextModel
from ..random_init import random_state_dict
from ..schedulers import B200EulerDiscreteScheduler, make_scheduler
from ..unet import B200UNet, unet_param_shapes
from ..vae import B200VAE, vae_param_shapes


def _load_safetensors(folder: str, stem: str) -> Dict[str, torch.Tensor]:
    from safetensors.torch import load_file

    for name in (f"{stem}.fp16.safetensors", f"{stem}.safetensors"):
        p = os.path.join(folder, name)
        if os.path.exists(p):
            return load_file(p)
    raise FileNotFoundError(f"no {stem}[.fp16].safetensors under {folder}")


def _resolve_local(ckpt_path: str) -> str:
    if os.path.isdir(ckpt_path):
        return ckpt_path
    raise FileNotFoundError(
        f"checkpoint '{ckpt_path}' is not a local folder; this build has no network access to the HF hub "
        "(models/stable_diffusion.py:110-152 would download it)")


class _ImageProcessor:
    """The part of diffusers' VaeImageProcessor the pipeline uses (`preprocess`, pipeline :238): tensors or PIL images
    -> float tensor (B,3,H,W) in [-1,1], H and W rounded down to a multiple of the VAE scale factor."""

    def __init__(self, vae_scale_factor: int, do_normalize: bool = True):
        self.f = vae_scale_factor
        self.do_normalize = do_normalize

    @staticmethod
    def get_crop_region(mask_image, width: int, height: int, pad: int = 0):
        """`mask_processor.get_crop_region` (pipeline :271; diffusers VaeImageProcessor): the rectangle (x1, y1, x2, y2) that
        holds every non-zero mask pixel, grown by `pad`, then expanded to the aspect ratio of the (width, height) the pipeline
        processes at, and shifted back inside the image."""
        import numpy as np

        if isinstance(mask_image, (list, tuple)):
            mask_image = mask_image[0]
        if not hasattr(mask_image, "convert"):
            raise ValueError("padding_mask_crop needs PIL images (the reference's get_crop_region works on a PIL mask)")
        mask = np.array(mask_image.convert("L"))
        h, w = mask.shape
        cols = np.nonzero(mask.any(axis=0))[0]
        rows = np.nonzero(mask.any(axis=1))[0]
        crop_left, crop_right = (int(cols[0]), int(w - 1 - cols[-1])) if len(cols) else (w, w)
        crop_top, crop_bottom = (int(rows[0]), int(h - 1 - rows[-1])) if len(rows) else (h, h)
        x1, y1 = int(max(crop_left - pad, 0)), int(max(crop_top - pad, 0))
        x2, y2 = int(min(w - crop_right + pad, w)), int(min(h - crop_bottom + pad, h))
        ratio_crop_region = (x2 - x1) / (y2 - y1)
        ratio_processing = width / height
        if ratio_crop_region > ratio_processing:
            desired_height = (x2 - x1) / ratio_processing
            diff = int(desired_height - (y2 - y1))
            y1 -= diff // 2
            y2 += diff - diff // 2
            if y2 >= h:
                d = y2 - h
                y2 -= d
                y1 -= d
            if y1 < 0:
                y2 -= y1
                y1 -= y1
            if y2 >= h:
                y2 = h
        else:
            desired_width = (y2 - y1) * ratio_processing
            diff = int(desired_width - (x2 - x1))
            x1 -= diff // 2
            x2 += diff - diff // 2
            if x2 >= w:
                d = x2 - w
                x2 -= d
                x1 -= d
            if x1 < 0:
                x2 -= x1
                x1 -= x1
            if x2 >= w:
                x2 = w
        return x1, y1, x2, y2

    @staticmethod
    def _resize_and_fill(im, width: int, height: int):
        """resize_mode="fill" (pipeline :272): keep the aspect ratio, centre, fill what is left with the stretched border."""
        from PIL import Image

        ratio, src_ratio = width / height, im.width / im.height
        src_w = width if ratio < src_ratio else im.width * height // im.height
        src_h = height if ratio >= src_ratio else im.height * width // im.width
        resized = im.resize((src_w, src_h), resample=Image.LANCZOS)
        res = Image.new(im.mode, (width, height))
        res.paste(resized, box=(width // 2 - src_w // 2, height // 2 - src_h // 2))
        if ratio < src_ratio:
            fill = height // 2 - src_h // 2
            if fill > 0:
                res.paste(resized.resize((width, fill), box=(0, 0, width, 0)), box=(0, 0))
                res.paste(resized.resize((width, fill), box=(0, resized.height, width, resized.height)), box=(0, fill + src_h))
        elif ratio > src_ratio:
            fill = width // 2 - src_w // 2
            if fill > 0:
                res.paste(resized.resize((fill, height), box=(0, 0, 0, height)), box=(0, 0))
                res.paste(resized.resize((fill, height), box=(resized.width, 0, resized.width, height)), box=(fill + src_w, 0))
        return res

    def _pil_resize(self, im, width, height, crops_coords, resize_mode):
        w, h = im.size
        w, h = (width or w) // self.f * self.f, (height or h) // self.f * self.f
        if crops_coords is not None:
            im = im.crop(crops_coords)
        if resize_mode == "fill":
            return self._resize_and_fill(im, w, h)
        if resize_mode != "default":
            raise ValueError(f"resize_mode {resize_mode} is not supported")
        return im.resize((w, h))

    def preprocess(self, image, height=None, width=None, crops_coords=None, resize_mode="default", **unused):
        import numpy as np

        if isinstance(image, torch.Tensor):
            if crops_coords is not None:
                raise ValueError("padding_mask_crop needs PIL images")
            x = image if image.dim() == 4 else image[None]
            return x  # tensors are taken as already in [-1, 1] (diffusers does the same for tensor input in [-1,1])
        imgs = image if isinstance(image, list) else [image]
        arrs = []
        for im in imgs:
            im = self._pil_resize(im.convert("RGB"), width, height, crops_coords, resize_mode)
            arrs.append(np.asarray(im, dtype=np.float32) / 255.0)
        x = torch.from_numpy(np.stack(arrs)).permute(0, 3, 1, 2)
        return 2.0 * x - 1.0 if self.do_normalize else x

    def preprocess_mask(self, mask, height=None, width=None, crops_coords=None, resize_mode="default"):
        """`mask_processor.preprocess` (pipeline :283-285; VaeImageProcessor(do_normalize=False, do_binarize=True,
        do_convert_grayscale=True)): one channel in {0, 1}, 1 = repaint, resized to (height, width)."""
        import numpy as np

        if isinstance(mask, torch.Tensor):
            if crops_coords is not None:
                raise ValueError("padding_mask_crop needs PIL images")
            x = mask.float()
            while x.dim() < 4:
                x = x[None]
            if x.shape[1] != 1:
                x = x.mean(dim=1, keepdim=True)
        else:
            masks = mask if isinstance(mask, list) else [mask]
            if crops_coords is not None or resize_mode != "default":
                masks = [self._pil_resize(mm.convert("L"), width, height, crops_coords, resize_mode) for mm in masks]
            x = torch.from_numpy(np.stack([np.asarray(mm.convert("L"), dtype=np.float32) / 255.0 for mm in masks]))[:, None]
        if height is not None and width is not None and tuple(x.shape[-2:]) != (height, width):
            x = torch.nn.functional.interpolate(x, size=(height, width))
        return (x >= 0.5).float()


def _is_hf_clip(enc) -> bool:
    return type(enc).__name__ in ("CLIPTextModel", "CLIPTextModelWithProjection") and hasattr(enc, "state_dict")


class SDModelWrapper:
    def __init__(self, ckpt_path: Optional[str] = None, vae=None, base=None, text_encoder=None, tokenizer=None,
                 scheduler=None, text_encoder_2=None, tokenizer_2=None, refiner=None, model_name: Optional[str] = None,
                 model_type: Optional[str] = None, device: str = "cuda"):
        """Same argument list as the reference (:41-55).  Components passed explicitly are used as they are;
        everything else is loaded from `ckpt_path` (default names as in the reference :79-88)."""
        self.type = model_type or "sd15"
        if self.type == "sd15":
            for a in ("text_encoder_2", "tokenizer_2"):
                if hasattr(self, a):
                    delattr(self, a)
        else:
            self.text_encoder_2 = text_encoder_2
            self.tokenizer_2 = tokenizer_2
        if model_name is not None and isinstance(model_name, str):
            ckpt_path = f"OnMoon/{self.type}_{model_name}"
        if ckpt_path is None and base is None:
            ckpt_path = "stabilityai/stable-diffusion-xl-base-1.0" if self.type == "sdxl" else "runwayml/stable-diffusion-v1-5"
        self.vae, self.base, self.text_encoder, self.tokenizer, self.scheduler = vae, base, text_encoder, tokenizer, scheduler
        if refiner is not None:
            self.refiner = refiner
        if base is None:
            self.load_hf_checkpoint(ckpt_path)
        if self.scheduler is None:
            self.scheduler = B200EulerDiscreteScheduler(**scheduler_config(self.type))
        self.device = torch.device("cpu")
        self.to(device)
        self.vae_scale_factor = 2 ** (len(self.vae.config.block_out_channels) - 1)
        self.lora_loader = None  # diffusers' LoraLoaderMixin is replaced by stablediffusion_b200.lora (merge at load)
        self.image_processor = _ImageProcessor(self.vae_scale_factor)
        self.mask_processor = _ImageProcessor(self.vae_scale_factor, do_normalize=False)
        self.scheduler_name = "euler"
        self.name = model_name or None
        self.path = ckpt_path

    # ---- construction without a checkpoint (benchmarks / tests: random-init weights of the named architecture) ----
    @classmethod
    def from_random(cls, model_type: str = "sd15", seed: int = 0, device: str = "cuda", unet_overrides=None,
                    vae_overrides=None, with_vae_encoder: bool = False):
        ucfg = unet_config(model_type, **(unet_overrides or {}))
        vcfg = vae_config(model_type, **(vae_overrides or {}))
        usd = random_state_dict(unet_param_shapes(ucfg), seed=seed)
        vsd = random_state_dict(vae_param_shapes(vcfg, decoder=True, encoder=with_vae_encoder), seed=seed + 1)
        base = B200UNet(ucfg, usd, device=device)
        vae = B200VAE(vcfg, vsd, device=device)
        sched = B200EulerDiscreteScheduler(**scheduler_config(model_type))
        return cls(vae=vae, base=base, scheduler=sched, model_type=model_type, device=device)

    def load_hf_checkpoint(self, ckpt_path):
        if hasattr(self, "path") and self.path == ckpt_path:
            return
        root = _resolve_local(ckpt_path)
        vcfg = load_json_config(os.path.join(root, "vae"), VAE_DEFAULTS)
        self.vae = B200VAE(vcfg, _load_safetensors(os.path.join(root, "vae"), "diffusion_pytorch_model"), device="cpu")
        ucfg = load_json_config(os.path.join(root, "unet"), UNET_DEFAULTS)
        self.base = B200UNet(ucfg, _load_safetensors(os.path.join(root, "unet"), "diffusion_pytorch_model"), device="cpu")
        from transformers import CLIPTextModel, CLIPTextModelWithProjection, CLIPTokenizer

        self.text_encoder = CLIPTextModel.from_pretrained(root, subfolder="text_encoder", torch_dtype=torch.bfloat16)
        self.tokenizer = CLIPTokenizer.from_pretrained(root, subfolder="tokenizer")
        scfg = load_json_config(os.path.join(root, "scheduler"), {}, "scheduler_config.json")
        self.scheduler = B200EulerDiscreteScheduler(**scfg)
        if hasattr(self, "text_encoder_2"):
            self.text_encoder_2 = CLIPTextModelWithProjection.from_pretrained(root, subfolder="text_encoder_2", torch_dtype=torch.bfloat16)
            self.tokenizer_2 = CLIPTokenizer.from_pretrained(root, subfolder="tokenizer_2")
        self.path = ckpt_path
        print(f"StableDiffusion model successfully loaded from '{ckpt_path}' checkpoint")

    def reload(self, ckpt_path=None, model_name=None, model_type=None):
        self.__init__(ckpt_path=ckpt_path, model_name=model_name, model_type=model_type, device=self.device)

    def to(self, device):
        self.vae.to(device)
        self.base.to(device)
        # text encoders: transformers modules become native B200 encoders the first time they reach the device
        # (SURVEY §8(f) rank 2); the torch module stays on the host as the master copy LoRA is merged into
        for name in ("text_encoder", "text_encoder_2"):
            enc = getattr(self, name, None)
            if enc is None:
                continue
            if torch.device(device).type == "cuda" and not isinstance(enc, B200CLIPTextModel) and _is_hf_clip(enc):
                setattr(self, name, B200CLIPTextModel.from_transformers(enc, device=device))
            else:
                enc.to(device)
        self.device = torch.device(device)

    def set_scheduler(self, scheduler_name):
        if hasattr(self, "scheduler_name") and self.scheduler_name == scheduler_name:
            return
        self.scheduler = make_scheduler(scheduler_name, self.scheduler.config)  # raises ValueError on unknown names (:224)
        self.scheduler_name = scheduler_name
        print(f"Scheduler has successfully changed to '{scheduler_name}'")

    # ---- LoRA (reference :230-335).  Adapters are merged into the packed weights, not run as extra GEMMs. ----
    def load_loras(self, loras, lora_dir: Optional[str] = None):
        if isinstance(loras, str):
            loras = {loras: 1.0}
        elif isinstance(loras, list):
            loras = {name: 1.0 for name in loras}
        current = [] if self.get_list_adapters() == {} else self.get_list_adapters()["base"]
        self.delete_adapters(current)
        for lora_name in list(loras.keys()):
            fname = f"{self.type}_{lora_name}.safetensors"  # hub naming of the reference (:242-245), resolved locally
            path = os.path.join(lora_dir or os.environ.get("B200SD_LORA_DIR", "."), fname)
            if not os.path.exists(path):
                raise FileNotFoundError(f"LoRA file {path} not found (no network: 'OnMoon/loras' cannot be downloaded)")
            self.load_lora_weights(path, adapter_name=lora_name)
        print(f"LoRA adapters has successfully changed to:\n{loras}")
        self.set_adapters(list(loras.keys()), list(loras.values()))

    def load_lora_weights(self, pretrained_model_name_or_path_or_dict: Union[str, Dict[str, torch.Tensor]],
                          adapter_name: Optional[str] = None, **kwargs):
        if isinstance(pretrained_model_name_or_path_or_dict, dict):
            sd = pretrained_model_name_or_path_or_dict
        else:
            from safetensors.torch import load_file

            p = pretrained_model_name_or_path_or_dict
            if os.path.isdir(p):
                p = os.path.join(p, "pytorch_lora_weights.safetensors")  # train_lora_pipeline.py:497-528 output name
            sd = load_file(p)
        modules = sorted({k.rsplit(".", 1)[0] for k in unet_param_shapes(self.base.config) if k.endswith(".weight")})
        parsed = parse_lora_state_dict(sd, unet_modules=modules)  # raises ValueError("Invalid LoRA checkpoint.") (:264-266)
        adapter_name = adapter_name or f"default_{len(self.base.peft_config)}"
        self.base.add_lora(adapter_name, parsed["unet"])
        for part, enc_name in (("text_encoder", "text_encoder"), ("text_encoder_2", "text_encoder_2")):
            enc = getattr(self, enc_name, None)
            if parsed[part] and enc is not None:
                self._merge_text_encoder_lora(enc, parsed[part], adapter_name)

    def _merge_text_encoder_lora(self, enc, layers, adapter_name):
        """Text-encoder LoRA (outside the hot path): kept as (module, delta) pairs; the encoder's weights are always
        RE-MERGED from an fp32 pristine copy (like the UNet's), so load / set / delete cycles do not accumulate rounding."""
        store = enc.__dict__.setdefault("_b200_lora", {})
        mods = dict(enc.named_modules())
        flat = {name.replace(".", "_"): name for name in mods}  # kohya spelling: lora_te_text_model_encoder_layers_0_...
        from ..lora import lora_delta

        ent, missing = [], []
        for path, e in layers.items():
            body = path
            for pfx in ("lora_te1_", "lora_te2_", "lora_te_"):
                if body.startswith(pfx):
                    body = body[len(pfx):]
                    break
            name = body if body in mods else flat.get(body)
            if name is None:
                missing.append(path)
                continue
            m = mods[name]
            if not hasattr(m, "_b200_pristine"):
                m._b200_pristine = m.weight.data.detach().float().clone()
            ent.append((m, lora_delta(e)))
        if missing:
            raise ValueError(f"LoRA adapter '{adapter_name}' targets {len(missing)} text-encoder modules that do not exist, "
                             f"e.g. {missing[:3]}")
        store[adapter_name] = dict(layers=ent, weight=0.0)
        enc.peft_config = {k: True for k in store}

    def _apply_text_encoder_weights(self, enc, names, weights):
        store = getattr(enc, "_b200_lora", None)
        if not store:
            return
        want = {n: (1.0 if w is None else float(w)) for n, w in zip(names, weights)}
        changed = False
        for name, st in store.items():
            new_w = want.get(name, 0.0)
            changed |= new_w != st["weight"]
            st["weight"] = new_w
        if not changed:
            return
        touched = {}
        for st in store.values():
            for m, delta in st["layers"]:
                acc = touched.setdefault(id(m), [m, m._b200_pristine.clone()])
                if st["weight"] != 0.0:
                    acc[1] += st["weight"] * delta.reshape(acc[1].shape).to(acc[1].device)
        for m, w in touched.values():
            m.weight.data.copy_(w.to(m.weight.device, m.weight.dtype))
        if isinstance(enc, B200CLIPTextModel):
            enc.refresh_weights()

    def set_adapters(self, adapter_names: Union[List[str], str], adapter_weights: Optional[List[float]] = None):
        names = [adapter_names] if isinstance(adapter_names, str) else list(adapter_names)
        weights = adapter_weights if adapter_weights is not None else [1.0] * len(names)
        self.base.set_adapters(names, weights)
        for enc_name in ("text_encoder", "text_encoder_2"):
            enc = getattr(self, enc_name, None)
            if enc is not None:
                self._apply_text_encoder_weights(enc, names, weights)

    def delete_adapters(self, adapter_names: Union[List[str], str]):
        if isinstance(adapter_names, str):
            adapter_names = [adapter_names]
        self.base.delete_adapters(adapter_names)
        for enc_name in ("text_encoder", "text_encoder_2"):
            enc = getattr(self, enc_name, None)
            store = getattr(enc, "_b200_lora", None) if enc is not None else None
            if store:
                self._apply_text_encoder_weights(enc, [], [])
                for n in adapter_names:
                    store.pop(n, None)
                enc.peft_config = {k: True for k in store}

    def get_list_adapters(self) -> Dict[str, List[str]]:
        out = {}
        if len(self.base.peft_config) > 0:
            out["base"] = list(self.base.peft_config.keys())
        for enc_name in ("text_encoder", "text_encoder_2"):
            enc = getattr(self, enc_name, None)
            if enc is not None and getattr(enc, "_b200_lora", None):
                out[enc_name] = list(enc._b200_lora.keys())
        return out
