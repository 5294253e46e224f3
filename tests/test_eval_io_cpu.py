"""Oracle restatements and host mirrors of the eval-side data formats against the reference's own outputs
(tests/golden/eval_io.npz, made by oracle/make_golden_eval.py from datasets/ray_utils.py, metrics.py, eval.py:201-203)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import nerfw_oracle as O                    # noqa: E402

G = np.load(os.path.join(ROOT, "tests", "golden", "eval_io.npz"))


def _case(name):
    return int(G[f"{name}/H"]), int(G[f"{name}/W"]), G[f"{name}/K"], torch.from_numpy(G[f"{name}/c2w"])


def test_oracle_rays_match_reference():
    for name in ("odd", "blender"):
        H, W, K, c2w = _case(name)
        assert torch.equal(O.ray_directions(H, W, K), torch.from_numpy(G[f"{name}/directions"]))
        rays = O.rays_of_view(H, W, K, c2w, 2.0, 6.0)
        ref = torch.from_numpy(G[f"{name}/rays"])
        assert rays.shape == ref.shape == (H * W, 8)
        assert (rays - ref).abs().max().item() <= 1e-6            # bmm vs mm accumulation order
        assert torch.equal(rays[:, [0, 1, 2, 6, 7]], ref[:, [0, 1, 2, 6, 7]])


def test_oracle_u8_and_psnr_match_reference():
    x, gt = torch.from_numpy(G["u8/x"]), torch.from_numpy(G["psnr/gt"])
    assert torch.equal(O.pack_rgb8(x), torch.from_numpy(G["u8/y"]))
    assert abs(float(O.psnr(x, gt)) - float(G["psnr/raw"])) <= 1e-6
    assert abs(float(O.psnr(gt, x.clamp(0, 1))) - float(G["psnr/clipped"])) <= 1e-6
    import pytest
    from nerf_fl_b200 import metrics as M                     # the product mirror has no CPU path
    with pytest.raises(RuntimeError, match="no CPU path"):
        M.psnr(x, gt)
