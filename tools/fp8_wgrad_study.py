"""CPU study (oracle only, no kernels): what would 8-bit saved images cost in gradient parity?

The training step moves ~4.6 GB per 1024 rays because the forward saves every layer's bf16 activation image and
dgrad writes every layer's bf16 dZ image for the wgrad kernel.  Storing both as e4m3 (per-128-row-tile scale for dZ)
would halve that.  This script swaps the oracle's bf16-operand field for one whose WEIGHT-gradient contraction
dW = dZ^T . A sees e4m3-rounded operands (the dgrad chain stays bf16), and reports the gradient-vector error against
the fp32 oracle next to the bf16-only model's -- the north_star bar is 1e-2 relative.

    python tools/fp8_wgrad_study.py [n_rays]
"""
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from nerf_fl_b200 import synthetic as syn          # noqa: E402
from oracle import nerfw_oracle as O                # noqa: E402

MODE = {"act": "bf16", "dz": "bf16"}


def q_e4m3(x, tile_rows=128, scaled=True):
    """round to e4m3 with one power-of-two scale per tile of rows (amax -> 256)"""
    if not scaled:
        return x.to(torch.float8_e4m3fn).float()
    n = x.shape[0]
    pad = (-n) % tile_rows
    xp = F.pad(x, (0, 0, 0, pad)).view(-1, tile_rows, x.shape[1])
    amax = xp.abs().amax(dim=(1, 2), keepdim=True).clamp_min(1e-30)
    s = torch.exp2(torch.floor(torch.log2(256.0 / amax)))
    y = ((xp * s).to(torch.float8_e4m3fn).float() / s).view(-1, x.shape[1])
    return y[:n]


def rnd(x, kind):
    if kind == "bf16":
        return x.to(torch.bfloat16).float()
    if kind == "e4m3":
        return q_e4m3(x, scaled=False)
    if kind == "e4m3s":
        return q_e4m3(x)
    if kind == "e5m2s":
        n = x.shape[0]
        pad = (-n) % 128
        xp = F.pad(x, (0, 0, 0, pad)).view(-1, 128, x.shape[1])
        amax = xp.abs().amax(dim=(1, 2), keepdim=True).clamp_min(1e-30)
        s = torch.exp2(torch.floor(torch.log2(16384.0 / amax)))
        return ((xp * s).to(torch.float8_e5m2).float() / s).view(-1, x.shape[1])[:n]
    if kind.startswith("e5m2g"):
        # ONE power-of-two scale per pass, from amax of the pass's d_raw (what a kernel can know before the chain starts):
        # amax -> 2^T
        T = int(kind[5:])
        import math
        sc = 2.0 ** (T - math.ceil(math.log2(max(MODE["amax"], 1e-30))))
        y = (x * sc).clamp(-57344.0, 57344.0).to(torch.float8_e5m2).float() / sc
        return y
    raise KeyError(kind)


class LinQ(torch.autograd.Function):
    """y = bf16(x) . bf16(W)^T ; backward: dx = bf16(dy) . bf16(W) ; dW = q(dy)^T . q(x)"""

    @staticmethod
    def forward(ctx, x, W):
        xb, Wb = x.to(torch.bfloat16).float(), W.to(torch.bfloat16).float()
        ctx.save_for_backward(xb, Wb)
        return xb @ Wb.t()

    @staticmethod
    def backward(ctx, dy):
        xb, Wb = ctx.saved_tensors
        dyb = dy.to(torch.bfloat16).float()
        dx = dyb @ Wb
        dW = rnd(dyb, MODE["dz"]).t() @ rnd(xb, MODE["act"])
        return dx, dW


def split_e5m2(x, terms):
    """x ~ sum of `terms` e5m2 numbers (each the rounding of the residual so far), under the pass's global scale"""
    tot = torch.zeros_like(x)
    for _ in range(terms):
        tot = tot + rnd(x - tot, MODE["dz"])
    return tot


def split_e4m3(x, terms):
    tot = torch.zeros_like(x)
    for _ in range(terms):
        tot = tot + (x - tot).to(torch.float8_e4m3fn).float()
    return tot


class HeadQ(torch.autograd.Function):
    """fp32 head (as in the kernels); its WEIGHT gradient contracts 8-bit images: dpre (e5m2, MODE['head_dp'] terms)
    against the head's input activation (e4m3, MODE['head_act'] terms)"""

    @staticmethod
    def forward(ctx, h, W, b):
        ctx.save_for_backward(h, W)
        return F.linear(h, W, b)

    @staticmethod
    def backward(ctx, dy):
        h, W = ctx.saved_tensors
        dh = dy @ W
        if MODE.get("head_dp", 0) == 0:
            dW = dy.t() @ h
        else:
            dW = split_e5m2(dy, MODE["head_dp"]).t() @ split_e4m3(h, MODE["head_act"])
        return dh, dW, dy.sum(0)


def head(p, name, h):
    return HeadQ.apply(h, p[name + ".weight"], p[name + ".bias"])


def field_q(p, xyz, dir_a, tt, output_transient, sigma_only=False):
    def lin(n, x, cols=None):
        W = p[n + ".weight"]
        return LinQ.apply(x, W if cols is None else W[:, cols[0]:cols[1]])

    h = xyz
    for i in range(8):
        n = f"xyz_encoding_{i + 1}.0"
        if i == 4:
            h = torch.cat([xyz, h], 1)
        h = torch.relu(lin(n, h) + p[n + ".bias"])
    sigma = F.softplus(head(p, "static_sigma.0", h))
    if sigma_only:
        return sigma
    feat = lin("xyz_encoding_final", h) + p["xyz_encoding_final.bias"]
    Wd = p["dir_encoding.0.weight"]
    hd = torch.relu(lin("dir_encoding.0", feat, (0, 256)) + F.linear(dir_a, Wd[:, 256:], p["dir_encoding.0.bias"]))
    out = [torch.sigmoid(head(p, "static_rgb.0", hd)), sigma]
    if output_transient:
        Wt = p["transient_encoding.0.weight"]
        ht = torch.relu(lin("transient_encoding.0", feat, (0, 256)) + F.linear(tt, Wt[:, 256:], p["transient_encoding.0.bias"]))
        for j in (2, 4, 6):
            n = f"transient_encoding.{j}"
            ht = torch.relu(lin(n, ht) + p[n + ".bias"])
        out += [torch.sigmoid(head(p, "transient_rgb.0", ht)), F.softplus(head(p, "transient_sigma.0", ht)),
                F.softplus(head(p, "transient_beta.0", ht))]
    res = torch.cat(out, 1)
    if res.requires_grad:
        res.register_hook(lambda g: MODE.__setitem__("amax", g.abs().max().item()))
    return res


def grads(n, scale, model):
    gen = torch.Generator().manual_seed(7)
    rays = syn.blender_rays(n, seed=3)
    ts = torch.randint(0, 100, (n,), generator=gen)
    tgt = torch.rand(n, 3, generator=gen)
    rng = {"perturb_rand": torch.rand(n, 64, generator=gen), "u": torch.rand(n, 64, generator=gen)}
    pc = syn.make_field_params(41, "coarse", weight_scale=scale)
    pf = syn.make_field_params(42, "fine", True, True, weight_scale=scale)
    ea, et = syn.make_embedding(31, 100, 48), syn.make_embedding(32, 100, 16)
    named = [(f"coarse.{k}", v) for k, v in pc.items()] + [(f"fine.{k}", v) for k, v in pf.items()] + [("emb_a", ea), ("emb_t", et)]
    for _, v in named:
        v.requires_grad_(True)
    orig = O.field_forward_bf16
    if model not in ("fp32", "bf16"):
        O.field_forward_bf16 = field_q
    try:
        out = O.render_rays(pc, pf, ea, et, rays, ts, n_samples=64, n_importance=64, perturb=1.0, noise_std=0.0,
                            white_back=True, encode_a=True, encode_t=True, beta_min=0.1, rng=rng,
                            bf16_operands=(model != "fp32"))
    finally:
        O.field_forward_bf16 = orig
    loss = sum(O.nerfw_loss(out, tgt).values())
    g = torch.autograd.grad(loss, [v for _, v in named])
    return {k: x.double() for (k, _), x in zip(named, g)}


def rel(a, b, keys):
    num = sum(((a[k] - b[k]).norm() ** 2).item() for k in keys)
    return (num / sum((b[k].norm() ** 2).item() for k in keys)) ** 0.5


def study_heads(n, scale):
    g32 = grads(n, scale, "fp32")
    keys = list(g32)
    MODE["act"], MODE["dz"] = "e4m3", "e5m2g10"
    print(f"heads through the 8-bit contraction too (n_rays {n}): dpre terms / activation terms -> error vs fp32")
    for hd, ha in ((0, 0), (1, 1), (3, 1), (1, 2), (3, 2)):
        MODE["head_dp"], MODE["head_act"] = hd, ha
        g = grads(n, scale, "q")
        ck = [k for k in keys if k.startswith("coarse.")]
        worst = sorted(((rel(g, g32, [k]), k) for k in keys if "sigma" in k or "rgb" in k or "beta" in k), reverse=True)[:3]
        print(f"  dp x{hd} act x{ha}: model {rel(g, g32, keys):.2e} coarse {rel(g, g32, ck):.2e} fine {rel(g, g32, [k for k in keys if k not in ck]):.2e}"
              f"  heads worst {[(round(r, 4), k) for r, k in worst]}")
    MODE["head_dp"] = 0


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[2] == "heads":
        torch.set_num_threads(os.cpu_count())
        study_heads(int(sys.argv[1]), 1.0)
        sys.exit(0)
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    torch.set_num_threads(os.cpu_count())
    for scale in (1.0,):
        g32 = grads(n, scale, "fp32")
        keys = list(g32)
        rows = []
        for act, dz in (("bf16", "bf16"), ("e4m3", "bf16"), ("bf16", "e4m3s"), ("e4m3", "e4m3s"), ("e4m3", "e5m2s"),
                        ("e4m3", "e5m2g0"), ("e4m3", "e5m2g6"), ("e4m3", "e5m2g10"), ("e4m3", "e5m2g14")):
            MODE["act"], MODE["dz"] = act, dz
            g = grads(n, scale, "q")
            worst = max(((rel(g, g32, [k]), k) for k in keys))
            rows.append((act, dz, rel(g, g32, keys), rel(g, g32, [k for k in keys if k.startswith("coarse.")]),
                         rel(g, g32, [k for k in keys if not k.startswith("coarse.")]), worst))
        print(f"n_rays {n}  weights x{scale}: gradient-vector error vs the fp32 oracle (bar 1e-2)")
        for act, dz, allr, cr, fr, worst in rows:
            print(f"  saved act {act:6s} dZ {dz:6s}: model {allr:.2e}  coarse {cr:.2e}  fine+emb {fr:.2e}  worst tensor {worst[0]:.2e} ({worst[1]})")
        # per tensor against the bf16-only model (the test suite's 3e-2 per-tensor bar is against that model)
        MODE["act"], MODE["dz"] = "bf16", "bf16"
        gb = grads(n, scale, "q")
        for dzk in ("e4m3s", "e5m2g6", "e5m2g10"):
            MODE["act"], MODE["dz"] = "e4m3", dzk
            g8 = grads(n, scale, "q")
            per = sorted(((rel(g8, gb, [k]), k) for k in keys), reverse=True)[:4]
            print(f"  e4m3/{dzk} vs bf16-only model: model {rel(g8, gb, keys):.2e}, worst tensors:", [(round(r, 4), k) for r, k in per])
