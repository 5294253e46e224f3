"""Generate tests/golden/eval_io.npz: rays, uint8 image packing and PSNR from the UNMODIFIED reference
(datasets/ray_utils.py, metrics.py, the eval.py:201-203 numpy lines).  Build container only:

    python oracle/make_golden_eval.py

kornia is not installed here.  ray_utils.py uses exactly one kornia function, create_meshgrid(H, W,
normalized_coordinates=False) (kornia==0.4.1, requirements.txt): a (1, H, W, 2) fp32 grid whose [..., 0] is the
column index x in [0, W-1] and [..., 1] the row index y in [0, H-1].  A stub module with that one function (and a
placeholder kornia.losses.ssim, unused here) is put in sys.modules so the reference files import unchanged;
the files are loaded by path so datasets/__init__.py (PIL, torchvision datasets) is not pulled in.
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"


def _stub_kornia():
    k = types.ModuleType("kornia")

    def create_meshgrid(height, width, normalized_coordinates=True, device=None):
        assert not normalized_coordinates
        xs = torch.linspace(0, width - 1, width, device=device, dtype=torch.float32)
        ys = torch.linspace(0, height - 1, height, device=device, dtype=torch.float32)
        gy, gx = torch.meshgrid(ys, xs, indexing="ij")
        return torch.stack([gx, gy], -1).unsqueeze(0)

    k.create_meshgrid = create_meshgrid
    losses = types.ModuleType("kornia.losses")
    losses.ssim = lambda *a, **kw: (_ for _ in ()).throw(NotImplementedError("ssim stub"))
    k.losses = losses
    sys.modules["kornia"] = k
    sys.modules["kornia.losses"] = losses


def _load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


_stub_kornia()
ray_utils = _load("ref_ray_utils", os.path.join(REF, "datasets", "ray_utils.py"))
metrics = _load("ref_metrics", os.path.join(REF, "metrics.py"))

out = {}
gen = torch.Generator().manual_seed(2024)


def pose(gen):
    q, _ = torch.linalg.qr(torch.randn(3, 3, generator=gen))
    return torch.cat([q, torch.randn(3, 1, generator=gen) * 3.0], 1).float()


# (1) odd-sized image, fx != fy, off-centre principal point
cases = {"odd": (7, 5, np.array([[61.5, 0, 2.25], [0, 58.0, 3.5], [0, 0, 1]])),
         # (2) blender.py:55-62 intrinsics at 40x40 (focal scaled from 800, fov 0.6911 rad, K[0,2] = w/2, K[1,2] = h/2)
         "blender": (40, 40, None)}
f = 0.5 * 800 / np.tan(0.5 * 0.6911) * 40 / 800
Kb = np.eye(3); Kb[0, 0] = Kb[1, 1] = f; Kb[0, 2] = 20; Kb[1, 2] = 20
cases["blender"] = (40, 40, Kb)
for name, (H, W, K) in cases.items():
    c2w = pose(gen)
    dirs = ray_utils.get_ray_directions(H, W, K)                       # (H, W, 3)
    rays_o, rays_d = ray_utils.get_rays(dirs, c2w)                     # (H*W, 3) each
    near, far = 2.0, 6.0
    rays = torch.cat([rays_o, rays_d, near * torch.ones_like(rays_o[:, :1]), far * torch.ones_like(rays_o[:, :1])], 1)  # blender.py:97-100
    out[f"{name}/H"], out[f"{name}/W"], out[f"{name}/K"], out[f"{name}/c2w"] = H, W, K, c2w
    out[f"{name}/directions"] = dirs.float()
    out[f"{name}/rays"] = rays.float()

# (3) eval.py:201-203
x = torch.rand(333, 3, generator=gen) * 1.4 - 0.2
x[:6, 0] = torch.tensor([0.0, 1.0, -1e-9, 1.0 + 1e-6, 254.999 / 255, 0.5])
x[6:262, 1] = torch.arange(256).float() / 255                            # the exact k/255 boundaries
img_pred = np.clip(x.view(-1, 3).cpu().numpy(), 0, 1)
out["u8/x"] = x
out["u8/y"] = (img_pred * 255).astype(np.uint8)

# (4) metrics.py:12, as train.py:166 calls it and as eval.py:208-210 does (on the clipped prediction)
gt = torch.rand(333, 3, generator=gen)
out["psnr/gt"] = gt
out["psnr/raw"] = metrics.psnr(x, gt)
out["psnr/mse_raw"] = metrics.mse(x, gt)
out["psnr/clipped"] = metrics.psnr(gt, torch.from_numpy(img_pred))

np.savez_compressed(os.path.join(ROOT, "tests", "golden", "eval_io.npz"),
                    **{k: (v.detach().numpy() if torch.is_tensor(v) else np.asarray(v)) for k, v in out.items()})
print("wrote eval_io.npz", {k: np.asarray(v.detach() if torch.is_tensor(v) else v).shape for k, v in out.items()})
