"""Training objective over the maps render_rays returns -- the role of the reference's ``losses.py``
(ColorLoss :4-16, NerfWLoss :19-50; NeRF-W eq. 13), which sits just outside the per-ray hot path
 and decides which outputs receive cotangents: rgb_coarse, rgb_fine, beta and
transient_sigmas.  ``NerfWLoss`` / ``ColorLoss`` / ``loss_dict`` keep the reference's names so train.py:80
can import them from here.  Both are one forward and one backward kernel (``csrc/objective.cu``: written with
torch ops the same reductions are ~40 launches of 2-3 us, a tenth of a 1024-ray training step).  There is no CPU
or ATen path: CPU tensors raise (the torch-op restatement the tests compare against lives in tests/host_models.py
and in the oracle).
"""
from __future__ import annotations

import torch
from torch import nn

__all__ = ["ColorLoss", "NerfWLoss", "loss_dict"]


class _NerfWLossFn(torch.autograd.Function):
    """(c_l, f_l, b_l, s_l, total) from one kernel; the backward is one kernel taking whichever of the five
    cotangents arrive (device scalars, graph-capturable: no value ever visits the host)."""

    @staticmethod
    def forward(ctx, rgb_coarse, rgb_fine, beta, transient_sigmas, targets, lambda_u, coef):
        from . import kernels as K
        tens = [t.contiguous() if t is not None else None for t in (rgb_coarse, rgb_fine, beta, transient_sigmas, targets)]
        terms = K.nerfw_loss_fwd(*tens, float(lambda_u), float(coef))
        ctx.tens, ctx.scal = tens, (float(lambda_u), float(coef))
        ctx.set_materialize_grads(False)
        return tuple(terms[i] for i in range(5))

    @staticmethod
    def backward(ctx, *g):
        from . import kernels as K
        g = [x.contiguous() if x is not None else None for x in g]
        d_c, d_f, d_b, d_s = K.nerfw_loss_bwd(*ctx.tens, *ctx.scal, g)
        return d_c, d_f, d_b, d_s, None, None, None


class LossTerms(dict):
    """The reference's dict of loss terms (train.py:166 sums its values); ``total`` is that sum, produced by
    the same kernel, so a caller that wants only the scalar saves the adds."""
    total = None


class NerfWLoss(nn.Module):
    def __init__(self, coef=1, lambda_u=0.01):
        super().__init__()
        self.coef, self.lambda_u = coef, lambda_u

    def forward(self, inputs, targets):
        if not targets.is_cuda:
            raise RuntimeError("nerf_fl_b200.NerfWLoss needs CUDA tensors: there is no CPU path")
        rgb_f, beta = inputs.get("rgb_fine"), inputs.get("beta")
        ts = inputs.get("transient_sigmas") if beta is not None else None
        c_l, f_l, b_l, s_l, total = _NerfWLossFn.apply(inputs["rgb_coarse"], rgb_f, beta, ts, targets, self.lambda_u, self.coef)
        out = LossTerms(c_l=c_l)
        if rgb_f is not None:
            out["f_l"] = f_l
        if beta is not None:
            out["b_l"], out["s_l"] = b_l, s_l
        out.total = total
        return out


class ColorLoss(nn.Module):
    """Plain-NeRF objective (losses.py:4-16): coef * (mse(rgb_coarse) + mse(rgb_fine)).  The same kernel pair as
    NerfWLoss without beta: its total is coef' * 0.5 * (mse_c + mse_f), so coef' = 2 * coef."""

    def __init__(self, coef=1):
        super().__init__()
        self.coef = coef

    def forward(self, inputs, targets):
        if not targets.is_cuda:
            raise RuntimeError("nerf_fl_b200.ColorLoss needs CUDA tensors: there is no CPU path")
        *_, total = _NerfWLossFn.apply(inputs["rgb_coarse"], inputs.get("rgb_fine"), None, None, targets, 0.0,
                                       2.0 * self.coef)
        return total


loss_dict = {"color": ColorLoss, "nerfw": NerfWLoss}
