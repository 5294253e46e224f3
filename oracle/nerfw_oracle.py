"""CPU oracle for the NeRF-W per-ray hot path.  TEST INFRASTRUCTURE ONLY.

This file is a from-scratch restatement (functional style, plain torch CPU ops, fp32) of the
algorithm in the reference's ``models/rendering.py`` and ``models/nerf.py``.  It exists to CHECK
the CUDA path.  Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import it; the product package ``nerf_fl_b200`` never does.

Parity status: PINNED.  The reference has no tests or golden vectors of its own, so
the pin is the reference itself: ``oracle/make_golden.py`` imports ``/root/reference`` (read-only,
in the build container), runs ``render_rays`` / ``NeRF`` / ``PosEmbedding`` / ``sample_pdf`` on
seeded inputs and commits inputs + outputs under ``tests/golden/``; ``tests/test_oracle_golden.py``
checks every function below against those vectors (bit-exact for indices, <=1e-6 for floats).

Every function cites the reference lines it follows (paths relative to /root/reference).
Parameters are passed as plain ``state_dict``-style dicts using the reference's key names
 so the same tensors drive oracle and product.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

__all__ = [
    "posenc", "field_forward", "composite", "sample_pdf", "coarse_depths", "render_rays",
    "nerfw_loss", "field_macs",
]


# ----------------------------------------------------------------------------------------------
# models/nerf.py:6-32  PosEmbedding
# ----------------------------------------------------------------------------------------------
def posenc(x: torch.Tensor, n_freqs: int) -> torch.Tensor:
    """[x, sin(2^0 x), cos(2^0 x), ..., sin(2^(L-1) x), cos(2^(L-1) x)]  (nerf.py:15,27-32).

    The reference builds ``freqs = 2**linspace(0, L-1, L)`` (exact powers of two, nerf.py:15) and
    multiplies before taking sin/cos (nerf.py:30), so the argument is an exact fp32 scaling.
    """
    parts = [x]
    for k in range(n_freqs):
        arg = x * float(2 ** k)
        parts.append(torch.sin(arg))
        parts.append(torch.cos(arg))
    return torch.cat(parts, dim=-1)


# ----------------------------------------------------------------------------------------------
# models/nerf.py:153-212  NeRF.forward
# ----------------------------------------------------------------------------------------------
def _lin(p, name, x):
    return F.linear(x, p[name + ".weight"], p[name + ".bias"])


def field_forward(p: dict, x: torch.Tensor, *, in_xyz: int = 63, in_dir_a: int = 27,
                  in_t: int = 16, sigma_only: bool = False, output_transient: bool = True,
                  D: int = 8, skips=(4,)) -> torch.Tensor:
    """8x256 trunk with the skip at layer index 4 (input FIRST in the concat, nerf.py:186-188),
    Softplus sigma head (nerf.py:190), linear 256->256 feature (nerf.py:194), ReLU(331->128) on
    cat[feat, dir(+a)] (nerf.py:195-196), Sigmoid rgb (nerf.py:197), transient branch on
    cat[feat, t] (nerf.py:203-207), output column order [rgb_s, sigma_s, rgb_t, sigma_t, beta]
    (nerf.py:198,209-212).
    """
    if sigma_only:
        xyz = x
    elif output_transient:
        xyz, dir_a, tt = torch.split(x, [in_xyz, in_dir_a, in_t], dim=-1)
    else:
        xyz, dir_a = torch.split(x, [in_xyz, in_dir_a], dim=-1)
    h = xyz
    for i in range(D):
        if i in skips:
            h = torch.cat([xyz, h], dim=1)
        h = torch.relu(_lin(p, f"xyz_encoding_{i + 1}.0", h))
    sigma = F.softplus(_lin(p, "static_sigma.0", h))
    if sigma_only:
        return sigma
    feat = _lin(p, "xyz_encoding_final", h)
    hd = torch.relu(_lin(p, "dir_encoding.0", torch.cat([feat, dir_a], dim=1)))
    rgb = torch.sigmoid(_lin(p, "static_rgb.0", hd))
    static = torch.cat([rgb, sigma], dim=1)
    if not output_transient:
        return static
    ht = torch.cat([feat, tt], dim=1)
    for j in (0, 2, 4, 6):
        ht = torch.relu(_lin(p, f"transient_encoding.{j}", ht))
    t_sigma = F.softplus(_lin(p, "transient_sigma.0", ht))
    t_rgb = torch.sigmoid(_lin(p, "transient_rgb.0", ht))
    t_beta = F.softplus(_lin(p, "transient_beta.0", ht))
    return torch.cat([static, t_rgb, t_sigma, t_beta], dim=1)


def field_macs(encode_a: bool, encode_t: bool, sigma_only: bool = False,
               in_xyz: int = 63, in_dir: int = 27, in_a: int = 48, in_t: int = 16,
               W: int = 256) -> int:
    """Multiply-accumulates per sample of ``field_forward``."""
    trunk = in_xyz * W + 3 * W * W + (W + in_xyz) * W + 3 * W * W
    if sigma_only:
        return trunk + W
    m = trunk + W + W * W + (W + in_dir + (in_a if encode_a else 0)) * (W // 2) + 3 * (W // 2)
    if encode_t:
        m += (W + in_t) * (W // 2) + 3 * (W // 2) ** 2 + 5 * (W // 2)
    return m


# ----------------------------------------------------------------------------------------------
# models/rendering.py:141-226  volume rendering inside inference()
# ----------------------------------------------------------------------------------------------
def _excl_cumprod_one_minus(alpha: torch.Tensor) -> torch.Tensor:
    """T_i = prod_{j<i}(1-alpha_j) as cumprod of [1, 1-a_0, ...][:-1]  (rendering.py:154-156)."""
    shifted = torch.cat([torch.ones_like(alpha[:, :1]), 1 - alpha], dim=-1)
    return torch.cumprod(shifted[:, :-1], dim=-1)


def composite(raw: torch.Tensor, z: torch.Tensor, *, typ: str, transient: bool,
              white_back: bool, test_time: bool, beta_min: float = 0.0,
              noise: torch.Tensor | None = None, noise_std: float = 0.0) -> dict:
    """raw (N,S,1|4|9), z (N,S) -> the result-dict entries of one inference() call.

    ``raw`` has 1 column for the test-time coarse pass (sigma only, rendering.py:103-111),
    4 for static only, 9 with the transient head.  ``noise`` is the standard-normal tensor of
    rendering.py:151 (pre-drawn by the caller so product and oracle can share it).
    """
    N, S = z.shape
    out = {}
    if raw.shape[-1] == 1:
        s_sigma = raw[..., 0]
    else:
        s_rgb, s_sigma = raw[..., :3], raw[..., 3]
    delta = torch.cat([z[:, 1:] - z[:, :-1], torch.full_like(z[:, :1], 1e2)], dim=-1)  # :142-144
    if transient:
        t_rgb, t_sigma, t_beta = raw[..., 4:7], raw[..., 7], raw[..., 8]
        a_s = 1 - torch.exp(-delta * s_sigma)                                           # :147
        a_t = 1 - torch.exp(-delta * t_sigma)                                           # :148
        alpha = 1 - torch.exp(-delta * (s_sigma + t_sigma))                             # :149
    else:
        nz = (noise if noise is not None else torch.zeros_like(s_sigma)) * noise_std    # :151
        alpha = 1 - torch.exp(-delta * torch.relu(s_sigma + nz))                        # :152
    T = _excl_cumprod_one_minus(alpha)
    w = alpha * T                                                                       # :162
    wsum = w.sum(dim=1)                                                                 # :163
    out[f"weights_{typ}"] = w
    out[f"opacity_{typ}"] = wsum
    if transient:
        out["transient_sigmas"] = t_sigma                                               # :168
    if test_time and typ == "coarse":
        return out                                                                      # :169-170
    if transient:
        w_s, w_t = a_s * T, a_t * T                                                     # :159-160
        s_map = (w_s.unsqueeze(-1) * s_rgb).sum(dim=1)
        if white_back:
            s_map = s_map + (1 - wsum.unsqueeze(-1))                                    # :176-177
        t_map = (w_t.unsqueeze(-1) * t_rgb).sum(dim=1)
        out["beta"] = (w_t * t_beta).sum(dim=1) + beta_min                              # :182-185
        out["_rgb_fine_static"] = s_map
        out["_rgb_fine_transient"] = t_map
        out["rgb_fine"] = s_map + t_map                                                 # :190
        if test_time:                                                                   # :192-217
            Ts = _excl_cumprod_one_minus(a_s)
            ws_only = a_s * Ts
            m = (ws_only.unsqueeze(-1) * s_rgb).sum(dim=1)
            if white_back:
                m = m + (1 - wsum.unsqueeze(-1))   # combined opacity on purpose (:203-204)
            out["rgb_fine_static"] = m
            out["depth_fine_static"] = (ws_only * z).sum(dim=1)
            Tt = _excl_cumprod_one_minus(a_t)
            wt_only = a_t * Tt
            out["rgb_fine_transient"] = (wt_only.unsqueeze(-1) * t_rgb).sum(dim=1)
            out["depth_fine_transient"] = (wt_only * z).sum(dim=1)
    else:
        m = (w.unsqueeze(-1) * s_rgb).sum(dim=1)                                        # :219-220
        if white_back:
            m = m + (1 - wsum.unsqueeze(-1))
        out[f"rgb_{typ}"] = m
    out[f"depth_{typ}"] = (w * z).sum(dim=1)                                            # :225
    return out


# ----------------------------------------------------------------------------------------------
# models/rendering.py:7-46  sample_pdf
# ----------------------------------------------------------------------------------------------
def sample_pdf(bins: torch.Tensor, weights: torch.Tensor, n_importance: int, *,
               det: bool = False, u: torch.Tensor | None = None, eps: float = 1e-5,
               return_aux: bool = False):
    """Inverse-CDF sampling.  bins (N,M+1), weights (N,M) -> (N,n_importance).

    ``u`` overrides the uniform draw (rendering.py:26-30) so callers can share RNG tensors.
    With ``return_aux`` also returns (cdf, u, inds) for the bit-exact index test.
    """
    N, M = weights.shape
    wt = weights + eps                                                                  # :20
    pdf = wt / wt.sum(dim=1, keepdim=True)                                              # :21
    cdf = torch.cat([torch.zeros_like(pdf[:, :1]), torch.cumsum(pdf, dim=-1)], dim=-1)  # :22-24
    if u is None:
        if det:
            u = torch.linspace(0, 1, n_importance).expand(N, n_importance)              # :27-28
        else:
            u = torch.rand(N, n_importance)                                             # :30
    u = u.contiguous()
    inds = torch.searchsorted(cdf, u, right=True)                                       # :33
    lo = (inds - 1).clamp_min(0)                                                        # :34
    hi = inds.clamp_max(M)                                                              # :35
    cdf_lo, cdf_hi = torch.gather(cdf, 1, lo), torch.gather(cdf, 1, hi)                 # :37-38
    bin_lo, bin_hi = torch.gather(bins, 1, lo), torch.gather(bins, 1, hi)               # :39
    denom = cdf_hi - cdf_lo
    denom = torch.where(denom < eps, torch.ones_like(denom), denom)                     # :41-42
    samples = bin_lo + (u - cdf_lo) / denom * (bin_hi - bin_lo)                         # :45
    if return_aux:
        return samples, cdf, u, inds
    return samples


# ----------------------------------------------------------------------------------------------
# models/rendering.py:244-259  coarse depths
# ----------------------------------------------------------------------------------------------
def coarse_depths(near: torch.Tensor, far: torch.Tensor, n_samples: int, use_disp: bool,
                  perturb: float, perturb_rand: torch.Tensor | None) -> torch.Tensor:
    steps = torch.linspace(0, 1, n_samples)
    if not use_disp:
        z = near * (1 - steps) + far * steps                                            # :246
    else:
        z = 1 / (1 / near * (1 - steps) + 1 / far * steps)                              # :248
    z = z.expand(near.shape[0], n_samples)
    if perturb > 0:                                                                     # :252-259
        mid = 0.5 * (z[:, :-1] + z[:, 1:])
        upper = torch.cat([mid, z[:, -1:]], dim=-1)
        lower = torch.cat([z[:, :1], mid], dim=-1)
        r = perturb_rand if perturb_rand is not None else torch.rand(z.shape)
        z = lower + (upper - lower) * (perturb * r)
    return z


# ----------------------------------------------------------------------------------------------
# models/rendering.py:49-62, 228-289  render_rays
# ----------------------------------------------------------------------------------------------
def render_rays(p_coarse: dict, p_fine: dict | None, emb_a: torch.Tensor | None,
                emb_t: torch.Tensor | None, rays: torch.Tensor, ts: torch.Tensor | None, *,
                n_samples: int = 64, use_disp: bool = False, perturb: float = 0.0,
                noise_std: float = 1.0, n_importance: int = 0, white_back: bool = False,
                test_time: bool = False, n_freq_xyz: int = 10, n_freq_dir: int = 4,
                encode_a: bool = False, encode_t: bool = False, beta_min: float = 0.03,
                rng: dict | None = None, a_embedded=None, t_embedded=None,
                output_transient: bool = True, view_dir=None, bf16_operands: bool = False,
                wgrad8: bool = False) -> dict:
    """Functional restatement of render_rays.  ``bf16_operands=True`` swaps the field for
    field_forward_bf16 (the operand-rounding model of the tensor-core path, bottom of this file), and ``wgrad8=True`` on top of it
    for field_forward_q8 (same forward; weight gradients through the 8-bit images the kernels keep).  ``rng`` may hold pre-drawn tensors
    ``perturb_rand`` (N,Ns) U[0,1), ``noise_coarse`` (N,Ns) N(0,1), ``u`` (N,Ni) U[0,1),
    ``noise_fine`` (N,Ns+Ni) N(0,1); anything missing is drawn here in the reference's order
    (SURVEY.md §3.3), so under one ``torch.manual_seed`` this matches the reference bit for bit.
    The ``chunk`` argument of the reference only slices the point batch (rendering.py:101-130)
    and does not change results, so it is not modelled.
    """
    rng = dict(rng or {})
    N = rays.shape[0]
    o, d = rays[:, 0:3], rays[:, 3:6]
    near, far = rays[:, 6:7], rays[:, 7:8]
    dir_e = posenc(view_dir if view_dir is not None else d, n_freq_dir)                 # :238
    in_xyz = 3 + 6 * n_freq_xyz
    in_dir = 3 + 6 * n_freq_dir

    if perturb > 0 and "perturb_rand" not in rng:
        rng["perturb_rand"] = torch.rand(N, n_samples)
    z = coarse_depths(near, far, n_samples, use_disp, perturb, rng.get("perturb_rand"))

    def run(p, typ, zv, transient, a_e, t_e, noise):
        S = zv.shape[1]
        xyz = (o.unsqueeze(1) + d.unsqueeze(1) * zv.unsqueeze(-1)).reshape(-1, 3)       # :261,273
        xe = posenc(xyz, n_freq_xyz)
        if bf16_operands:
            so = typ == "coarse" and test_time
            da = dir_e if a_e is None else torch.cat([dir_e, a_e], 1)
            fld = field_forward_q8 if wgrad8 else field_forward_bf16
            raw = fld(p, xe, None if so else da.repeat_interleave(S, dim=0),
                      t_e.repeat_interleave(S, dim=0) if transient else None, transient, sigma_only=so)
        elif typ == "coarse" and test_time:                                             # :103-111
            raw = field_forward(p, xe, sigma_only=True)
        else:                                                                           # :112-139
            cols = [xe, dir_e.repeat_interleave(S, dim=0)]
            in_dir_a = in_dir
            if a_e is not None:
                cols.append(a_e.repeat_interleave(S, dim=0))
                in_dir_a += a_e.shape[1]
            if transient:
                cols.append(t_e.repeat_interleave(S, dim=0))
            raw = field_forward(p, torch.cat(cols, dim=1), in_xyz=in_xyz, in_dir_a=in_dir_a,
                                in_t=(t_e.shape[1] if transient else 0),
                                output_transient=transient)
        raw = raw.reshape(N, S, -1)
        return composite(raw, zv, typ=typ, transient=transient, white_back=white_back,
                         test_time=test_time, beta_min=beta_min, noise=noise,
                         noise_std=noise_std)

    if "noise_coarse" not in rng:
        rng["noise_coarse"] = torch.randn(N, n_samples)                                 # :151
    res = run(p_coarse, "coarse", z, False, None, None, rng["noise_coarse"])
    if n_importance > 0:
        mid = 0.5 * (z[:, :-1] + z[:, 1:])                                              # :268
        z_new = sample_pdf(mid, res["weights_coarse"][:, 1:-1].detach(), n_importance,
                           det=(perturb == 0),
                           u=(rng.get("u") if perturb > 0 else None))                   # :269-270
        z = torch.sort(torch.cat([z, z_new], dim=-1), dim=-1)[0]                        # :272
        a_e = t_e = None
        if encode_a:                                                                    # :276-280
            a_e = a_embedded if a_embedded is not None else emb_a[ts]
        transient = bool(output_transient and encode_t)                                 # :281
        if transient:                                                                   # :282-286
            t_e = t_embedded if t_embedded is not None else emb_t[ts]
        noise_f = None
        if not transient:
            if "noise_fine" not in rng:
                rng["noise_fine"] = torch.randn(N, n_samples + n_importance)
            noise_f = rng["noise_fine"]
        res.update(run(p_fine, "fine", z, transient, a_e, t_e, noise_f))
    return res


# ----------------------------------------------------------------------------------------------
# losses.py:35-50  NerfWLoss (defines which outputs receive cotangents)
# ----------------------------------------------------------------------------------------------
def nerfw_loss(res: dict, targets: torch.Tensor, coef: float = 1.0, lambda_u: float = 0.01):
    terms = {"c_l": 0.5 * ((res["rgb_coarse"] - targets) ** 2).mean()}
    if "rgb_fine" in res:
        if "beta" not in res:
            terms["f_l"] = 0.5 * ((res["rgb_fine"] - targets) ** 2).mean()
        else:
            terms["f_l"] = ((res["rgb_fine"] - targets) ** 2
                            / (2 * res["beta"].unsqueeze(1) ** 2)).mean()
            terms["b_l"] = 3 + torch.log(res["beta"]).mean()
            terms["s_l"] = lambda_u * res["transient_sigmas"].mean()
    return {k: coef * v for k, v in terms.items()}


# ----------------------------------------------------------------------------------------------
# Operand-rounding model of the product's tensor-core path (test infrastructure, like the rest of
# this file).  Same algorithm as field_forward (models/nerf.py:153-212), but every GEMM operand
# that the CUDA kernels feed to tcgen05 -- activations and the matching weight columns -- is
# rounded to bf16 first (fp32 accumulate); per-ray columns (dir / appearance / transient inputs)
# and the 1|3|5-wide heads stay fp32, as in the kernels.  Rounding is straight-through for autograd.
# Needed because ReLU gates of pre-activations within ~2^-9 of zero differ between an fp32 and a
# bf16-operand forward, which makes per-sample gradients of the two differ by a few percent no
# matter how the kernel is written; against THIS model the kernels must agree tightly.
# ----------------------------------------------------------------------------------------------
class _RoundBf16(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        return x.to(torch.bfloat16).float()

    @staticmethod
    def backward(ctx, g):
        return g


def field_forward_bf16(p: dict, xyz: torch.Tensor, dir_a: torch.Tensor | None, tt: torch.Tensor | None,
                       output_transient: bool, sigma_only: bool = False) -> torch.Tensor:
    r = _RoundBf16.apply

    def lin(n, x):
        return F.linear(r(x), r(p[n + ".weight"]), p[n + ".bias"])

    h = xyz
    for i in range(8):
        if i == 4:
            h = torch.cat([xyz, h], 1)
        h = torch.relu(lin(f"xyz_encoding_{i + 1}.0", h))
    sigma = F.softplus(_lin(p, "static_sigma.0", h))
    if sigma_only:
        return sigma
    feat = lin("xyz_encoding_final", h)
    Wd = p["dir_encoding.0.weight"]
    hd = torch.relu(F.linear(r(feat), r(Wd[:, :256])) + F.linear(dir_a, Wd[:, 256:], p["dir_encoding.0.bias"]))
    out = [torch.sigmoid(_lin(p, "static_rgb.0", hd)), sigma]
    if output_transient:
        Wt = p["transient_encoding.0.weight"]
        ht = torch.relu(F.linear(r(feat), r(Wt[:, :256])) + F.linear(tt, Wt[:, 256:], p["transient_encoding.0.bias"]))
        for j in (2, 4, 6):
            ht = torch.relu(lin(f"transient_encoding.{j}", ht))
        out += [torch.sigmoid(_lin(p, "transient_rgb.0", ht)), F.softplus(_lin(p, "transient_sigma.0", ht)),
                F.softplus(_lin(p, "transient_beta.0", ht))]
    return torch.cat(out, 1)


# ----------------------------------------------------------------------------------------------
# Operand-rounding model of the 8-bit weight-gradient path (nerf_fl_b200/csrc/field.cuh "8-bit images").
# The forward is field_forward_bf16's.  In the backward the cotangent chain stays bf16 (dx = bf16(dy) . bf16(W)),
# but every WEIGHT / BIAS gradient of a tensor-core layer contracts the two images the kernels keep in HBM:
#   dW = e5m2(s dy)^T . e4m3(bf16 x) / s,   db = sum e5m2(s dy) / s,
# s = one power of two per backward pass, 2^(DZ_TOP - ceil(log2 amax|d_raw|)).  The fp32 heads keep their fp32
# forward and dgrad; their weight gradient contracts a three-term e5m2 expansion of s dpre with the e4m3 image of the
# head's input plus the e4m3 image of 16 x what that left over.  Per-ray operands (view direction, embeddings, the biases folded into them) never
# pass through an image: plain fp32 autograd.  Test infrastructure, like everything else in this file: the kernels
# must agree with THIS model tightly, and this model's distance from the fp32 functions above is what the north-star
# tolerance has to cover.
# ----------------------------------------------------------------------------------------------
DZ_TOP = 10
LO_SCALE = 16.0       # the residual image of a head's input is stored times 16 (field.cuh kLoScale)


def _e4m3(x: torch.Tensor) -> torch.Tensor:          # cvt.rn.satfinite.e4m3x2.f32
    return x.clamp(-448.0, 448.0).to(torch.float8_e4m3fn).float()


def _e5m2(x: torch.Tensor) -> torch.Tensor:          # cvt.rn.satfinite.e5m2x2.f32
    return x.clamp(-57344.0, 57344.0).to(torch.float8_e5m2).float()


def _expand(x: torch.Tensor, rnd, terms: int) -> torch.Tensor:
    """sum of `terms` 8-bit numbers, each the rounding of what the earlier ones left over"""
    tot = torch.zeros_like(x)
    for _ in range(terms):
        tot = tot + rnd(x - tot)
    return tot


class DzScale:
    """the pass's scale, set by a hook on the field's output cotangent (field_bwd.cu dz_amax_kernel)"""

    def __init__(self):
        self.s = 1.0

    def set_from(self, d_raw: torch.Tensor) -> None:
        amax = d_raw.abs().max().item()
        if amax == 0.0 or not math.isfinite(amax):
            self.s = 1.0
            return
        m, e = math.frexp(amax)                      # amax = m 2^e, 0.5 <= m < 1: ceil(log2 amax) = e, or e - 1 at a power of two
        self.s = 2.0 ** (DZ_TOP - (e - 1 if m == 0.5 else e))


class _Lin8(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, W, b, st):
        xb, Wb = x.to(torch.bfloat16).float(), W.to(torch.bfloat16).float()
        ctx.save_for_backward(xb, Wb)
        ctx.st, ctx.has_b = st, b is not None
        y = xb @ Wb.t()
        return y + b if b is not None else y

    @staticmethod
    def backward(ctx, dy):
        xb, Wb = ctx.saved_tensors
        s = ctx.st.s
        dx = dy.to(torch.bfloat16).float() @ Wb
        q = _e5m2(dy * s) / s
        return dx, q.t() @ _e4m3(xb), (q.sum(0) if ctx.has_b else None), None


class _Head8(torch.autograd.Function):
    @staticmethod
    def forward(ctx, h, W, b, st):
        ctx.save_for_backward(h, W)
        ctx.st = st
        return F.linear(h, W, b)

    @staticmethod
    def backward(ctx, dy):
        h, W = ctx.saved_tensors
        s = ctx.st.s
        hi = _e4m3(h)
        dW = (_expand(dy * s, _e5m2, 3) / s).t() @ (hi + _e4m3((h - hi) * LO_SCALE) / LO_SCALE)
        return dy @ W, dW, dy.sum(0), None


def field_forward_q8(p: dict, xyz: torch.Tensor, dir_a: torch.Tensor | None, tt: torch.Tensor | None,
                     output_transient: bool, sigma_only: bool = False) -> torch.Tensor:
    st = DzScale()

    def lin(n, x, cols=None, bias=True):
        W = p[n + ".weight"]
        return _Lin8.apply(x, W if cols is None else W[:, cols[0]:cols[1]], p[n + ".bias"] if bias else None, st)

    def head(n, h):
        return _Head8.apply(h, p[n + ".weight"], p[n + ".bias"], st)

    h = xyz
    for i in range(8):
        if i == 4:
            h = torch.cat([xyz, h], 1)
        h = torch.relu(lin(f"xyz_encoding_{i + 1}.0", h))
    sigma = F.softplus(head("static_sigma.0", h))
    if sigma_only:
        res = sigma
    else:
        feat = lin("xyz_encoding_final", h)
        Wd = p["dir_encoding.0.weight"]
        hd = torch.relu(lin("dir_encoding.0", feat, (0, 256), bias=False) + F.linear(dir_a, Wd[:, 256:], p["dir_encoding.0.bias"]))
        out = [torch.sigmoid(head("static_rgb.0", hd)), sigma]
        if output_transient:
            Wt = p["transient_encoding.0.weight"]
            ht = torch.relu(lin("transient_encoding.0", feat, (0, 256), bias=False) +
                            F.linear(tt, Wt[:, 256:], p["transient_encoding.0.bias"]))
            for j in (2, 4, 6):
                ht = torch.relu(lin(f"transient_encoding.{j}", ht))
            out += [torch.sigmoid(head("transient_rgb.0", ht)), F.softplus(head("transient_sigma.0", ht)),
                    F.softplus(head("transient_beta.0", ht))]
        res = torch.cat(out, 1)
    if res.requires_grad:
        res.register_hook(st.set_from)
    return res


# ---- data formats either side of render_rays: rays, uint8 images, PSNR -------------------------
def ray_directions(H: int, W: int, K) -> torch.Tensor:
    """datasets/ray_utils.py:5-26: (H, W, 3) camera-space directions; i = column, j = row, no +0.5 centring."""
    j, i = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing="ij")
    fx, fy, cx, cy = float(K[0][0]), float(K[1][1]), float(K[0][2]), float(K[1][2])
    return torch.stack([(i - cx) / fx, -(j - cy) / fy, -torch.ones_like(i)], -1)


def rays_of_view(H: int, W: int, K, c2w: torch.Tensor, near: float, far: float) -> torch.Tensor:
    """ray_utils.py:29-55 + blender.py:97-100: (H*W, 8) rows [rays_o, rays_d, near, far], row-major pixels."""
    d = ray_directions(H, W, K).reshape(-1, 3) @ c2w[:, :3].T            # :46
    d = d / torch.norm(d, dim=-1, keepdim=True)                          # :48
    o = c2w[:, 3].expand(d.shape)                                        # :50
    one = torch.ones_like(d[:, :1])
    return torch.cat([o, d, near * one, far * one], 1)


def pack_rgb8(x: torch.Tensor) -> torch.Tensor:
    """eval.py:201-203: (np.clip(x, 0, 1) * 255).astype(np.uint8) -- fp32 product, truncation."""
    return (x.float().clamp(0, 1) * 255).to(torch.uint8)


def psnr(pred: torch.Tensor, gt: torch.Tensor) -> torch.Tensor:
    """metrics.py:4-13."""
    return -10 * torch.log10(torch.mean((pred - gt) ** 2))
