"""The eval loop around render_rays (eval.py:76-106, 186-210 of the reference), kept on the device.

``batched_inference`` keeps the reference's name and arguments; the difference is where the results
live: the reference moves EVERY key of every chunk to the host (eval.py:105-106 -- 0.33 GB of
``weights_fine`` per 800x800 image), this one concatenates only ``keys`` and leaves them on the GPU.
``render_image`` is the whole per-image body of eval.py:186-205: rays generated on the device from the
pose, rendered in chunks, ``rgb_fine`` clipped and packed to uint8 on the device
(8f-4), optional PSNR against a ground-truth image (metrics.py:12) -- 3 B/pixel cross PCIe.
"""
from __future__ import annotations

import torch

from . import kernels as K
from .parallel import dist_info, gather_rays, shard_range
from .ray_utils import image_rays
from .rendering import render_rays

__all__ = ["batched_inference", "render_image"]


@torch.no_grad()
def batched_inference(models, embeddings, rays, ts, N_samples, N_importance, use_disp, chunk, white_back,
                      keys=("rgb_fine", "depth_fine"), **kwargs):
    """eval.py:76-106.  keys=None keeps every key render_rays returns (the reference's behaviour, on the device)."""
    B = rays.shape[0]
    results = {}
    for i in range(0, B, chunk):
        r = render_rays(models, embeddings, rays[i:i + chunk], ts[i:i + chunk] if ts is not None else None, N_samples,
                        use_disp, 0, 0, N_importance, chunk, white_back, test_time=True, **kwargs)
        for k, v in r.items():
            if keys is None or k in keys:
                results.setdefault(k, []).append(v)
    return {k: torch.cat(v, 0) for k, v in results.items()}


@torch.no_grad()
def render_image(models, embeddings, H, W, K_intr, c2w, ts_value, N_samples, N_importance, use_disp=False,
                 chunk=32 * 1024, white_back=False, near=2.0, far=6.0, rgbs=None, device=None, **kwargs):
    """One view: -> {'img_u8': (H,W,3) uint8, 'rgb': (H*W,3) fp32, 'depth': (H*W,), ['psnr': 0-dim]} on the device.

    With torch.distributed initialised each rank renders a contiguous pixel block and the blocks are gathered
    (SURVEY.md 8e); every rank returns the full image."""
    dev = torch.device(device) if device is not None else next(models["coarse"].parameters()).device
    rank, world = dist_info()
    lo, hi = shard_range(H * W, rank, world)
    typ = "fine" if N_importance > 0 and "fine" in models else "coarse"
    parts = []
    for first in range(lo, hi, chunk):
        n = min(chunk, hi - first)
        rays = image_rays(H, W, K_intr, c2w, near, far, first, n, dev)
        ts = torch.full((n,), int(ts_value), dtype=torch.long, device=dev) if ts_value is not None else None
        r = render_rays(models, embeddings, rays, ts, N_samples, use_disp, 0, 0, N_importance, chunk, white_back,
                        test_time=True, **kwargs)
        parts.append(torch.cat([r[f"rgb_{typ}"], r[f"depth_{typ}"][:, None]], 1))
    local = torch.cat(parts, 0) if parts else torch.zeros(0, 4, device=dev)
    full = gather_rays(local, H * W)
    rgb = full[:, :3].contiguous()
    out = {"rgb": rgb, "depth": full[:, 3], "img_u8": K.pack_rgb8(rgb).view(H, W, 3)}
    if rgbs is not None:
        out["psnr"] = K.mse_psnr(rgb, rgbs.to(dev).reshape(-1, 3), clip01=True)[1]      # eval.py:207-210
    return out
