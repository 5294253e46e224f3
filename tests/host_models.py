"""Torch-op restatements used by the tests only (never imported by the product package):
  * adam_update_torch -- torch.optim.Adam's update on FlatTrainState's flat buffers, injected as ``host_update`` so the
    sharding / all-reduce plumbing can run at world_size 2 over gloo on CPU tensors;
  * nerfw_terms -- losses.py:35-50 written with torch ops, the autograd reference for csrc/objective.cu."""
import math

import torch


def adam_update_torch(state, world: int, zero_grad: bool) -> None:
    b1, b2 = state.betas
    g = state.flat_g / world + state.weight_decay * state.flat_p
    state.exp_avg.mul_(b1).add_(g, alpha=1 - b1)
    state.exp_avg_sq.mul_(b2).addcmul_(g, g, value=1 - b2)
    bc1, bc2 = 1 - b1 ** state.step_count, 1 - b2 ** state.step_count
    state.flat_p.addcdiv_(state.exp_avg, state.exp_avg_sq.sqrt() / math.sqrt(bc2) + state.eps, value=-state.lr / bc1)
    if zero_grad:
        state.flat_g.zero_()


def _sq_err(results, key, targets):
    diff = results[key] - targets
    return diff * diff


def nerfw_terms(results: dict, targets: torch.Tensor, lambda_u: float = 0.01) -> dict:
    """c_l (:37); with a fine pass f_l = 0.5 mse (:40) or the beta-weighted error (:42-43) with b_l = 3 + mean(log beta)
    (:44) and s_l = lambda_u * mean(transient_sigmas) (:45)."""
    terms = {"c_l": _sq_err(results, "rgb_coarse", targets).mean() * 0.5}
    if "rgb_fine" not in results:
        return terms
    err = _sq_err(results, "rgb_fine", targets)
    beta = results.get("beta")
    if beta is None:
        terms["f_l"] = err.mean() * 0.5
        return terms
    terms["f_l"] = (err / (2.0 * beta[:, None] ** 2)).mean()
    terms["b_l"] = beta.log().mean() + 3.0
    terms["s_l"] = results["transient_sigmas"].mean() * lambda_u
    return terms


def color_loss(results: dict, targets: torch.Tensor, coef: float = 1.0) -> torch.Tensor:
    """losses.py:4-16."""
    total = _sq_err(results, "rgb_coarse", targets).mean()
    if "rgb_fine" in results:
        total = total + _sq_err(results, "rgb_fine", targets).mean()
    return total * coef
