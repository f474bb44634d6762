"""Output stage of the serving handler, on the device.

Mirrors `convert_pt_to_numpy` of the reference (runpod-worker/handler_logic.py:21-29): the pipeline's image tensor
([-1, 1], NCHW) becomes a list of uint8 HWC arrays.  The reference does it image by image with five eager torch
ops and a blocking `.cpu()` each; here ONE kernel (`B200SD_OP_IMAGE_U8`) converts the whole batch, and ONE copy
into pinned host memory brings 3 bytes per pixel back instead of 6 (fp16) — the arithmetic is rounded exactly like
torch's eager ops on the tensor's dtype, so the bytes are the reference's bytes (tests/test_output_stage.py).

No CPU fallback: a tensor that is not on an sm_100 device raises.
"""
from __future__ import annotations

from typing import Dict, List, Tuple

import numpy as np
import torch

from .. import lowering as L


class OutputStage:
    """Caches one single-op program + pinned host buffer per (shape, dtype, device)."""

    def __init__(self):
        self._cache: Dict[Tuple, dict] = {}

    def _state(self, images: torch.Tensor) -> dict:
        key = (tuple(images.shape), images.dtype, images.device.index)
        st = self._cache.get(key)
        if st is None:
            B, C, H, W = images.shape
            b = L.Builder(images.device)
            st = {
                "x": b.empty([B, C, H * W], images.dtype),
                "out": torch.empty((B, H * W, C), dtype=torch.uint8, device=images.device),
                "host": torch.empty((B, H, W, C), dtype=torch.uint8).pin_memory(),
            }
            b.image_u8(st["x"], st["out"])
            st["prog"] = b.prog
            self._cache[key] = st
        return st

    def __call__(self, images: torch.Tensor) -> List[np.ndarray]:
        if not isinstance(images, torch.Tensor) or images.dim() != 4:
            raise ValueError("convert_pt_to_numpy expects an image tensor [B, C, H, W]")
        if not images.is_cuda:
            raise RuntimeError("the B200 output stage needs the image tensor on an sm_100 device (no CPU fallback)")
        if images.dtype not in (torch.bfloat16, torch.float16, torch.float32):
            raise ValueError(f"unsupported image dtype {images.dtype}")
        st = self._state(images)
        B, C, H, W = images.shape
        stream = torch.cuda.current_stream(images.device)
        st["x"].copy_(images.reshape(B, C, H * W))
        st["prog"].run(stream.cuda_stream)
        st["host"].copy_(st["out"].view(B, H, W, C), non_blocking=True)
        stream.synchronize()
        host = st["host"].numpy()
        return [host[i].copy() for i in range(B)]


_default_stage = OutputStage()


def convert_pt_to_numpy(images: torch.Tensor) -> List[np.ndarray]:
    """Same name, argument and result as runpod-worker/handler_logic.py:21-29."""
    return _default_stage(images)
