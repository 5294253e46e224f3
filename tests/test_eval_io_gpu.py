"""(f3) on-device ray generation, (f4) uint8 output pack, psnr -- CUDA kernels through the C-ABI against the
reference's own outputs (tests/golden/eval_io.npz), the oracle, and size-independent properties at full image size;
render_image (rays made on the device) against batched_inference on host-made rays."""
import os
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import nerfw_oracle as O                    # noqa: E402

G = np.load(os.path.join(ROOT, "tests", "golden", "eval_io.npz"))
DEV = torch.device("cuda:0")


def test_gen_rays_matches_reference_golden():
    from nerf_fl_b200.ray_utils import image_rays
    for name in ("odd", "blender"):
        H, W, K, c2w = int(G[f"{name}/H"]), int(G[f"{name}/W"]), G[f"{name}/K"], torch.from_numpy(G[f"{name}/c2w"])
        ref = torch.from_numpy(G[f"{name}/rays"])
        got = image_rays(H, W, K, c2w, 2.0, 6.0, device=DEV).cpu()
        assert got.shape == ref.shape
        assert torch.equal(got[:, [0, 1, 2, 6, 7]], ref[:, [0, 1, 2, 6, 7]])          # origin, near, far: copies
        assert (got[:, 3:6] - ref[:, 3:6]).abs().max().item() <= 1e-6                 # fp32 3-term dot + normalise
        # ragged pixel ranges (chunks that start and end mid-row) are slices of the same image
        for first, n in ((0, 1), (3, 11), (H * W - 4, 4), (W - 1, W + 2), (H * W, 0)):
            part = image_rays(H, W, K, c2w, 2.0, 6.0, first, n, DEV).cpu()
            assert torch.equal(part, got[first:first + n])


def test_gen_rays_full_image_properties():
    from nerf_fl_b200.ray_utils import image_rays
    H = W = 800
    f = 0.5 * W / np.tan(0.5 * 0.6911)
    K = np.array([[f, 0, W / 2], [0, f, H / 2], [0, 0, 1]])
    q, _ = torch.linalg.qr(torch.randn(3, 3, generator=torch.Generator().manual_seed(3)))
    c2w = torch.cat([q, torch.tensor([[0.5], [-4.0], [1.0]])], 1).float()
    rays = image_rays(H, W, K, c2w, 2.0, 6.0, device=DEV)
    want = O.rays_of_view(H, W, K, c2w, 2.0, 6.0).to(DEV)
    assert (rays - want).abs().max().item() <= 1e-6
    assert (rays[:, 3:6].norm(dim=-1) - 1).abs().max().item() <= 2e-7
    # the centre pixel looks down the camera's -z axis; rotating back recovers the pixel grid
    centre = rays[(H // 2) * W + W // 2, 3:6].cpu()
    assert (centre + c2w[:, 2]).abs().max().item() <= 1e-6
    cam = rays[:, 3:6].cpu() @ c2w[:, :3]
    ij = torch.stack([cam[:, 0] / -cam[:, 2] * f + W / 2, cam[:, 1] / cam[:, 2] * f + H / 2], 1)
    jj, ii = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    assert (ij - torch.stack([ii, jj], -1).reshape(-1, 2)).abs().max().item() <= 2e-3


def test_gen_rays_rejects_bad_ranges():
    from nerf_fl_b200.ray_utils import image_rays
    K = np.eye(3)
    with pytest.raises(RuntimeError, match="outside"):
        image_rays(4, 4, K, torch.eye(4)[:3], 2.0, 6.0, 10, 7, DEV)
    with pytest.raises(ValueError):
        image_rays(4, 4, K, torch.eye(4), 2.0, 6.0, device=DEV)


def test_pack_rgb8_bit_exact():
    from nerf_fl_b200 import kernels as K
    x = torch.from_numpy(G["u8/x"])
    assert torch.equal(K.pack_rgb8(x.to(DEV)).cpu(), torch.from_numpy(G["u8/y"]))
    g = torch.Generator().manual_seed(8)
    for n in (1, 2, 3, 5, 800 * 800 * 3):                          # tails that are not a multiple of 4
        big = torch.rand(n, generator=g) * 1.5 - 0.25
        got = K.pack_rgb8(big.to(DEV)).cpu()
        assert got.dtype == torch.uint8 and torch.equal(got, O.pack_rgb8(big))
    assert K.pack_rgb8(torch.zeros(0, 3, device=DEV)).shape == (0, 3)


def test_psnr_matches_reference_and_is_deterministic():
    from nerf_fl_b200 import metrics as M
    x, gt = torch.from_numpy(G["u8/x"]).to(DEV), torch.from_numpy(G["psnr/gt"]).to(DEV)
    assert abs(float(M.psnr(x, gt)) - float(G["psnr/raw"])) <= 1e-4
    assert abs(float(M.mse(x, gt)) - float(G["psnr/mse_raw"])) <= 1e-7
    from nerf_fl_b200 import kernels as K
    assert abs(float(K.mse_psnr(x, gt, clip01=True)[1]) - float(G["psnr/clipped"])) <= 1e-4
    g = torch.Generator().manual_seed(4)
    a, b = torch.rand(800 * 800, 3, generator=g).to(DEV), torch.rand(800 * 800, 3, generator=g).to(DEV)
    r1, r2 = K.mse_psnr(a, b), K.mse_psnr(a, b)
    assert torch.equal(r1, r2)
    assert abs(float(r1[0]) - float(((a.double() - b.double()) ** 2).mean())) <= 1e-6
    # a loss call between two metric calls shares the scratch and must not disturb it
    from nerf_fl_b200.objective import NerfWLoss
    NerfWLoss()({"rgb_coarse": a[:100]}, b[:100])
    assert torch.equal(K.mse_psnr(a, b), r1)


def test_render_image_equals_batched_inference_on_host_rays():
    from nerf_fl_b200 import synthetic as syn
    from nerf_fl_b200.inference import batched_inference, render_image
    from nerf_fl_b200.nerf import NeRF, PosEmbedding
    H, W = 48, 40
    f = 0.5 * W / np.tan(0.5 * 0.6911)
    K = np.array([[f, 0, W / 2], [0, f, H / 2], [0, 0, 1]])
    c2w = torch.tensor([[1.0, 0, 0, 0.1], [0, 1, 0, -0.2], [0, 0, 1, 4.0]])
    models = {}
    for i, (typ, a, t) in enumerate((("coarse", False, False), ("fine", True, True))):
        m = NeRF(typ, encode_appearance=a, encode_transient=t, beta_min=0.1)
        m.load_state_dict(syn.make_field_params(50 + i, typ, a, t))
        models[typ] = m.to(DEV)
    emb = {"xyz": PosEmbedding(9, 10), "dir": PosEmbedding(3, 4),
           "a": torch.nn.Embedding.from_pretrained(syn.make_embedding(5, 10, 48)).to(DEV),
           "t": torch.nn.Embedding.from_pretrained(syn.make_embedding(6, 10, 16)).to(DEV)}
    rays = O.rays_of_view(H, W, K, c2w, 2.0, 6.0).to(DEV)
    ts = torch.full((H * W,), 3, dtype=torch.long, device=DEV)
    ref = batched_inference(models, emb, rays, ts, 16, 16, False, 700, True)
    assert set(ref) == {"rgb_fine", "depth_fine"}
    gt = torch.rand(H * W, 3, generator=torch.Generator().manual_seed(1))
    out = render_image(models, emb, H, W, K, c2w, 3, 16, 16, chunk=700, white_back=True, rgbs=gt)
    assert out["img_u8"].shape == (H, W, 3) and out["img_u8"].dtype == torch.uint8
    assert (out["rgb"] - ref["rgb_fine"]).abs().max().item() <= 1e-4          # rays differ by <= 1e-6
    assert (out["depth"] - ref["depth_fine"]).abs().max().item() <= 1e-3
    assert torch.equal(out["img_u8"].reshape(-1, 3).cpu(), O.pack_rgb8(out["rgb"].cpu()))
    assert abs(float(out["psnr"]) - float(O.psnr(out["rgb"].cpu().clamp(0, 1), gt))) <= 1e-4
    every = batched_inference(models, emb, rays[:64], ts[:64], 16, 16, False, 700, True, keys=None)
    assert {"weights_fine", "weights_coarse", "beta"} <= set(every)      # test_time: the coarse pass is sigma-only
