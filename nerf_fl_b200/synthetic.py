"""Synthetic Blender-/Phototourism-shaped inputs (no datasets exist on the box).

Restates the ray construction of the reference's ``datasets/blender.py:55-66`` and
``datasets/ray_utils.py:18-53`` (pixel grid without +0.5, dirs = [(i-W/2)/f, -(j-H/2)/f, -1],
rays_d normalised, near 2 / far 6) on seeded look-at poses; SURVEY.md §8(d) fixes the constants.
Also a seeded parameter factory producing ``state_dict``s with the reference's key names
 so oracle, reference and product can be driven by identical weights.
"""
from __future__ import annotations

import math

import torch

FOV_X = 0.6911112070083618  # nerf_synthetic camera_angle_x (lego); any fixed value is fine


def _look_at_poses(n_views: int, radius: float, gen: torch.Generator) -> torch.Tensor:
    az = torch.rand(n_views, generator=gen) * (2 * math.pi)
    el = (torch.rand(n_views, generator=gen) * 0.5 + 0.1) * (math.pi / 2)
    eye = torch.stack([radius * torch.cos(el) * torch.cos(az),
                       radius * torch.cos(el) * torch.sin(az),
                       radius * torch.sin(el)], dim=-1)
    fwd = -eye / eye.norm(dim=-1, keepdim=True)          # camera looks at the origin
    up = torch.tensor([0.0, 0.0, 1.0]).expand_as(fwd)
    right = torch.linalg.cross(fwd, up)
    right = right / right.norm(dim=-1, keepdim=True)
    cam_up = torch.linalg.cross(right, fwd)
    # OpenGL-style c2w: columns [right, up, -forward | eye]
    R = torch.stack([right, cam_up, -fwd], dim=-1)
    return torch.cat([R, eye.unsqueeze(-1)], dim=-1)      # (V,3,4)


def blender_rays(n_rays: int, H: int = 400, W: int = 400, n_views: int = 100, seed: int = 1,
                 phototourism: bool = False) -> torch.Tensor:
    """(n_rays, 8) fp32 [o, d, near, far]: a seeded random draw of pixels over views."""
    gen = torch.Generator().manual_seed(seed)
    poses = _look_at_poses(n_views, 4.0, gen)
    focal = 0.5 * W / math.tan(0.5 * FOV_X)
    pick = torch.randint(0, n_views * H * W, (n_rays,), generator=gen)
    v = pick // (H * W)
    j = (pick % (H * W)) // W
    i = pick % W
    dirs = torch.stack([(i - W / 2) / focal, -(j - H / 2) / focal, -torch.ones(n_rays)], dim=-1)
    R, t = poses[v, :, :3], poses[v, :, 3]
    d = torch.einsum("nij,nj->ni", R, dirs.float())
    d = d / d.norm(dim=-1, keepdim=True)
    if phototourism:
        near_v = torch.rand(n_views, generator=gen) * 0.9 + 0.1
        far_v = torch.rand(n_views, generator=gen) * 2.0 + 3.0
        near, far = near_v[v], far_v[v]
    else:
        near, far = torch.full((n_rays,), 2.0), torch.full((n_rays,), 6.0)
    return torch.cat([t, d, near[:, None], far[:, None]], dim=-1).float().contiguous()


def view_camera(H: int, W: int, view: int = 0, seed: int = 1):
    """(K (3,3) float64 ndarray-like tensor, c2w (3,4) fp32) of the view image_rays(H, W, view, seed) draws: blender.py:55-62
    intrinsics (focal from FOV_X, principal point at W/2, H/2) and the seeded look-at pose."""
    gen = torch.Generator().manual_seed(seed)
    pose = _look_at_poses(max(view + 1, 1), 4.0, gen)[view]
    focal = 0.5 * W / math.tan(0.5 * FOV_X)
    K = torch.tensor([[focal, 0.0, W / 2], [0.0, focal, H / 2], [0.0, 0.0, 1.0]], dtype=torch.float64)
    return K, pose.float().contiguous()


def image_rays(H: int, W: int, view: int = 0, seed: int = 1) -> torch.Tensor:
    """All H*W rays of one view in row-major pixel order (the eval.py:186-194 shape)."""
    gen = torch.Generator().manual_seed(seed)
    pose = _look_at_poses(max(view + 1, 1), 4.0, gen)[view]
    focal = 0.5 * W / math.tan(0.5 * FOV_X)
    jj, ii = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    dirs = torch.stack([(ii - W / 2) / focal, -(jj - H / 2) / focal, -torch.ones(H, W)], dim=-1)
    d = dirs.reshape(-1, 3).float() @ pose[:, :3].T
    d = d / d.norm(dim=-1, keepdim=True)
    o = pose[:, 3].expand_as(d)
    n = H * W
    return torch.cat([o, d, torch.full((n, 1), 2.0), torch.full((n, 1), 6.0)], dim=-1).contiguous()


def field_param_shapes(typ: str, encode_a: bool = False, encode_t: bool = False,
                       in_xyz: int = 63, in_dir: int = 27, in_a: int = 48, in_t: int = 16,
                       W: int = 256, D: int = 8, skips=(4,)) -> dict:
    """name -> (out, in) of every Linear, in the reference's registration order (nerf.py:121-151).
    'coarse' silently disables appearance/transient (nerf.py:115-117)."""
    if typ == "coarse":
        encode_a = encode_t = False
    shapes = {}
    for i in range(D):
        fan_in = in_xyz if i == 0 else (W + in_xyz if i in skips else W)
        shapes[f"xyz_encoding_{i + 1}.0"] = (W, fan_in)
    shapes["xyz_encoding_final"] = (W, W)
    shapes["dir_encoding.0"] = (W // 2, W + in_dir + (in_a if encode_a else 0))
    shapes["static_sigma.0"] = (1, W)
    shapes["static_rgb.0"] = (3, W // 2)
    if encode_t:
        shapes["transient_encoding.0"] = (W // 2, W + in_t)
        for j in (2, 4, 6):
            shapes[f"transient_encoding.{j}"] = (W // 2, W // 2)
        shapes["transient_sigma.0"] = (1, W // 2)
        shapes["transient_rgb.0"] = (3, W // 2)
        shapes["transient_beta.0"] = (1, W // 2)
    return shapes


def make_field_params(seed: int, typ: str, encode_a: bool = False, encode_t: bool = False,
                      weight_scale: float = 1.0, **kw) -> dict:
    """Seeded U(-1/sqrt(fan_in), 1/sqrt(fan_in)) weights and biases (the bounds of nn.Linear's
    default init), weights optionally scaled (the "x3" parity variant of SURVEY.md §8c)."""
    gen = torch.Generator().manual_seed(seed)
    sd = {}
    for name, (fo, fi) in field_param_shapes(typ, encode_a, encode_t, **kw).items():
        bound = 1.0 / math.sqrt(fi)
        sd[name + ".weight"] = ((torch.rand(fo, fi, generator=gen) * 2 - 1) * bound
                                * weight_scale).contiguous()
        sd[name + ".bias"] = ((torch.rand(fo, generator=gen) * 2 - 1) * bound).contiguous()
    return sd


def make_embedding(seed: int, n_vocab: int, dim: int) -> torch.Tensor:
    """N(0,1) table, nn.Embedding's default init."""
    gen = torch.Generator().manual_seed(seed)
    return torch.randn(n_vocab, dim, generator=gen)
