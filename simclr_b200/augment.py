"""SimCLR's two-view augmentation on the GPU.

The reference augments on the host: ``SimCLRTransforms(strength, size)`` (dataset.py:42-50) is the ``transform`` of
the torchvision CIFAR dataset, so every image goes through PIL twice inside the data-loader workers (main.py:158-173)
before ``view0.to(local_rank)`` (main.py:112).  ``SimCLRTransformsGPU`` takes the raw uint8 batch on the device and
makes both views in one kernel launch (``ntx_simclr_augment``): same operations, same parameter distributions, PIL's
8-bit arithmetic.

    transform = SimCLRTransformsGPU(strength=0.5, size=32, seed=7)
    view0, view1 = transform(images_uint8)        # (B, 32, 32, 3) uint8 CUDA -> two (B, 3, 32, 32) float32 in [0, 1]
"""
from __future__ import annotations

import torch

from . import _cabi


class SimCLRTransformsGPU:
    """
    :param strength: colour distortion strength ``s`` of dataset.py:19-25 (brightness / contrast / saturation 0.8 s, hue 0.2 s).
    :param size: side of the output crops (dataset.py:43 default 32).
    :param seed: seed of the Philox streams; every call advances an internal offset, so calls differ and a run repeats.
    """

    def __init__(self, strength: float = 0.5, size: int = 32, seed: int = 0) -> None:
        assert strength > 0.
        self.strength, self.size, self.seed = strength, size, seed
        self.calls = 0

    def __call__(self, images: torch.Tensor, return_params: bool = False):
        if images.device.type != "cuda":
            raise RuntimeError("SimCLRTransformsGPU has no CPU path: pass the uint8 batch as a CUDA tensor")
        if images.dim() != 4 or images.shape[3] != 3 or images.dtype != torch.uint8:
            raise TypeError(f"expected a (B, H, W, 3) uint8 batch, got {tuple(images.shape)} {images.dtype}")
        images = images.contiguous()
        b = images.shape[0]
        views = torch.empty((2, b, 3, self.size, self.size), dtype=torch.float32, device=images.device)
        params = torch.empty((b, 2, 16), dtype=torch.float32, device=images.device) if return_params else None
        with torch.cuda.device(images.device):
            _cabi.Backend().simclr_augment(images, self.strength, self.seed, self.calls, self.size, views[0], views[1], params)
        self.calls += 1
        return (views[0], views[1], params) if return_params else (views[0], views[1])
