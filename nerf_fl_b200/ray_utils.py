"""Rays of a view for the render path: ``image_rays`` makes the (n, 8) rows
[rays_o, rays_d, near, far] of a pixel range of one view with one kernel on the device (csrc/raygen.cu restates
datasets/ray_utils.py:5-55 + blender.py:97-100), so an eval loop ships a 48-byte pose instead of 32 B/ray.
Dataset classes that build rays on the host once per image keep the reference's own ``get_ray_directions`` /
``get_rays`` (datasets/ray_utils.py); nothing on this path calls them.  There is no CPU path."""
from __future__ import annotations

import torch

from . import kernels as K_

__all__ = ["image_rays"]


def image_rays(H: int, W: int, K, c2w, near: float, far: float, first_pixel: int = 0, n_rays: int | None = None,
               device=None, out: torch.Tensor | None = None) -> torch.Tensor:
    """(n, 8) rows [rays_o, rays_d, near, far] (blender.py:97-100) of pixels [first_pixel, first_pixel + n) of the view
    with pose c2w, generated on the GPU."""
    return K_.gen_rays(H, W, K, c2w, near, far, first_pixel, n_rays, device, out)
