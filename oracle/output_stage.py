"""CPU oracle of the serving output stage — TEST INFRASTRUCTURE (see oracle/__init__.py for who may import it).

Restates `convert_pt_to_numpy` (runpod-worker/handler_logic.py:21-29) with the same torch ops in the same order on
the CPU; the reference function is five lines of eager torch, so this oracle IS pinned: torch's elementwise kernels
round every intermediate to the tensor's dtype, on CPU exactly as on CUDA, and `.to(torch.uint8)` truncates.
"""
from typing import List

import numpy as np
import torch


def convert_pt_to_numpy(images: torch.Tensor) -> List[np.ndarray]:
    images = images.detach().cpu()
    np_images = []
    for idx in range(len(images)):
        img = (images[idx] / 2 + 0.5).clamp(0, 1)                      # handler_logic.py:24
        img = (img.permute(1, 2, 0) * 255).to(torch.uint8).numpy()     # handler_logic.py:25
        np_images.append(img)
    return np_images
