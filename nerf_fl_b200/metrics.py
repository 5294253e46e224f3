"""metrics.py:4-13 of the reference (mse, psnr): the reduction is one deterministic kernel (csrc/objective.cu,
nwb_mse_psnr).  ``valid_mask`` selects elements first (a gather, no arithmetic).  Only ``reduction='mean'`` exists
here -- the reference never asks for anything else (eval.py:208, train.py:172,192) -- and the result carries no
gradient (the reference calls psnr under no_grad / on detached maps).  CUDA tensors only."""
from __future__ import annotations

from . import kernels as K

__all__ = ["mse", "psnr"]


def _pair(image_pred, image_gt, valid_mask, reduction):
    if reduction != "mean":
        raise NotImplementedError("nerf_fl_b200.metrics: only reduction='mean' is implemented")
    if not (image_pred.is_cuda and image_gt.is_cuda):
        raise RuntimeError("nerf_fl_b200.metrics needs CUDA tensors: there is no CPU path")
    pred, gt = image_pred.detach(), image_gt.detach()
    if valid_mask is not None:
        pred, gt = pred[valid_mask], gt[valid_mask]
    return K.mse_psnr(pred, gt)


def mse(image_pred, image_gt, valid_mask=None, reduction="mean"):
    return _pair(image_pred, image_gt, valid_mask, reduction)[0]


def psnr(image_pred, image_gt, valid_mask=None, reduction="mean"):
    return _pair(image_pred, image_gt, valid_mask, reduction)[1]
