"""GPU parity of the CUDA-core kernels (a1, a2, a5, a6) and of the tcgen05 tile engine, through
the C ABI, against the oracle and the committed golden vectors."""
import pytest
import torch

from nerf_fl_b200 import kernels as K
from nerf_fl_b200 import layout, synthetic as syn
from oracle import nerfw_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"


def maxerr(a, b):
    return (a.double().cpu() - b.double().cpu()).abs().max().item()


# ---- (a1) coarse depths ------------------------------------------------------------------------
@pytest.mark.parametrize("use_disp", [False, True])
@pytest.mark.parametrize("perturb", [0.0, 1.0, 0.5])
@pytest.mark.parametrize("ns", [64, 33, 128])
def test_sample_coarse_bit_exact(use_disp, perturb, ns):
    rays = syn.blender_rays(257, seed=7, phototourism=True)
    g = torch.Generator().manual_seed(1)
    pr = torch.rand(257, ns, generator=g)
    want = O.coarse_depths(rays[:, 6:7], rays[:, 7:8], ns, use_disp, perturb, pr)
    got = K.sample_coarse(rays.to(DEV), ns, use_disp, perturb, pr.to(DEV))
    if use_disp:
        assert maxerr(got, want) <= 5e-7 * want.abs().max().item()   # division rounding only
    else:
        assert torch.equal(got.cpu(), want)


def test_sample_coarse_philox_in_range():
    rays = syn.blender_rays(4096, seed=7).to(DEV)
    z0 = K.sample_coarse(rays, 64, False, 0.0)
    z1 = K.sample_coarse(rays, 64, False, 1.0, None, seed=123)
    z2 = K.sample_coarse(rays, 64, False, 1.0, None, seed=124)
    step = 4.0 / 63
    assert ((z1 - z0).abs() <= step / 2 + 1e-5).all() and not torch.equal(z1, z2)
    assert (z1[:, 1:] >= z1[:, :-1]).all()
    jit = ((z1 - z0) / step)[:, 1:-1]
    assert abs(jit.mean().item()) < 0.01 and abs(jit.std().item() - (1 / 12) ** 0.5) < 0.01


def test_philox_streams_of_the_two_draw_sites_are_independent():
    """the coarse stratified jitter (rendering.py:257) and the inverse-CDF uniforms (rendering.py:30) are
    drawn from the same seed; they must come from disjoint Philox counter blocks (draw ids 0 / 1), not be the same
    numbers.  Both are recovered from the kernels' outputs: z = lower + (upper-lower) u for the jitter, and with
    uniform weights the inverse CDF is linear in u."""
    n, Ns = 512, 64
    rays = syn.blender_rays(n, seed=7).to(DEV)
    z0 = K.sample_coarse(rays, Ns, False, 0.0)
    z1 = K.sample_coarse(rays, Ns, False, 1.0, None, seed=99)
    step = (z0[:, 1] - z0[:, 0])[:, None]
    u_c = ((z1 - (z0 - step / 2)) / step)[:, 1:-1]
    w = torch.ones(n, Ns, device=DEV)
    _, z_new, _ = K.sample_pdf_merge(z0, w, Ns, det=False, seed=99, debug=True)
    lo, hi = (z0[:, 0] + z0[:, 1]) / 2, (z0[:, -2] + z0[:, -1]) / 2
    u_f = ((z_new - lo[:, None]) / (hi - lo)[:, None])[:, 1:-1]
    assert u_c.min() >= -1e-4 and u_c.max() <= 1 + 1e-4 and u_f.min() >= -1e-4 and u_f.max() <= 1 + 1e-4
    assert (u_c - u_f).abs().mean().item() > 0.25                      # independent uniforms: E|a - b| = 1/3
    corr = torch.corrcoef(torch.stack([u_c.flatten(), u_f.flatten()]))[0, 1].item()
    assert abs(corr) < 0.03, corr


# ---- (a2) PosEmbedding -------------------------------------------------------------------------
@pytest.mark.parametrize("L", [4, 10, 15])
def test_posenc_golden(golden, L):
    g = golden("posenc")
    got = K.posenc(g["x"].to(DEV), L)
    assert got.shape == g[f"L{L}"].shape
    assert maxerr(got, g[f"L{L}"]) <= 4e-6      # libdevice sinf/cosf vs libm: a few ulp near 1


def test_posenc_empty_and_large():
    assert K.posenc(torch.zeros(0, 3, device=DEV), 10).shape == (0, 63)
    x = (torch.rand(100003, 3, generator=torch.Generator().manual_seed(2)) * 8 - 4)
    assert maxerr(K.posenc(x.to(DEV), 10), O.posenc(x, 10)) <= 4e-6


# ---- (a6) sample_pdf + merge -------------------------------------------------------------------
def _pdf_inputs(g):
    # golden bins/weights are free-standing; rebuild a (z, w) pair whose midpoints/inner weights
    # reproduce them is not possible in general, so the KAT drives the kernel with z = edges that
    # the kernel turns into midpoints: use the oracle on the kernel's own midpoints instead.
    N, Ns = 40, 64
    gen = torch.Generator().manual_seed(5)
    z = torch.sort(torch.rand(N, Ns, generator=gen) * 4 + 2, dim=-1)[0]
    w = torch.rand(N, Ns, generator=gen) ** 4
    w[3] = 0.0
    w[5, 10:50] = 0.0
    u = torch.rand(N, 64, generator=gen)
    u[7, -1] = 1.0
    u[8, 0] = 0.0
    return z, w, u


@pytest.mark.parametrize("det", [False, True])
def test_sample_pdf_indices_bit_exact_given_cdf(det):
    z, w, u = _pdf_inputs(None)
    mid = 0.5 * (z[:, :-1] + z[:, 1:])
    s, cdf, uu, inds = O.sample_pdf(mid, w[:, 1:-1], 64, det=det, u=None if det else u, return_aux=True)
    zo, zn, ii = K.sample_pdf_merge(z.to(DEV), w.to(DEV), 64, det, None if det else u.to(DEV),
                                    cdf=cdf.to(DEV), debug=True)
    assert torch.equal(ii.cpu(), inds)                       # bit-exact bin indices
    assert torch.equal(zn.cpu(), s)                          # same fp32 op order -> identical samples
    want = torch.sort(torch.cat([z, s], -1), -1)[0]
    assert torch.equal(zo.cpu(), want)
    assert inds.max().item() == 63 or det


# sort widths: 32 (padded), 64, 128 and 256 keys in registers, 512 through the shared-memory network
@pytest.mark.parametrize("ns,ni", [(64, 64), (128, 128), (64, 128), (128, 64), (33, 17), (16, 5), (64, 200), (64, 300), (8, 1)])
def test_sample_pdf_merge_own_cdf(ns, ni):
    gen = torch.Generator().manual_seed(9)
    N = 301
    z = torch.sort(torch.rand(N, ns, generator=gen) * 4 + 2, dim=-1)[0]
    w = torch.rand(N, ns, generator=gen) ** 2 + 0.01
    u = torch.rand(N, ni, generator=gen)
    mid = 0.5 * (z[:, :-1] + z[:, 1:])
    s = O.sample_pdf(mid, w[:, 1:-1], ni, det=False, u=u)
    want = torch.sort(torch.cat([z, s], -1), -1)[0]
    zo, zn, ii = K.sample_pdf_merge(z.to(DEV), w.to(DEV), ni, False, u.to(DEV), debug=True)
    assert (zo[:, 1:] >= zo[:, :-1]).all()                   # sortedness
    # the merge itself is exact: the output is sort(cat) of the coarse depths and the kernel's own new samples
    assert torch.equal(zo.cpu(), torch.sort(torch.cat([z, zn.cpu()], -1), -1)[0])
    # the in-kernel CDF is a warp scan, ATen's a sequential sum: cdf differs by ~1 ulp, which the
    # inverse-CDF lerp turns into <= ulp/pdf_bin of a bin width
    assert maxerr(zo, want) <= 1e-4, maxerr(zo, want)
    # adversarial weights (near-empty bins): order / range / multiset-size properties only
    w2 = torch.rand(N, ns, generator=gen) ** 8
    z2 = K.sample_pdf_merge(z.to(DEV), w2.to(DEV), ni, False, u.to(DEV))
    assert (z2[:, 1:] >= z2[:, :-1]).all() and torch.isfinite(z2).all()
    # Philox path: in range, sorted, right shape
    zr = K.sample_pdf_merge(z.to(DEV), w.to(DEV), ni, False, None, seed=3)
    assert zr.shape == (N, ns + ni) and (zr[:, 1:] >= zr[:, :-1]).all()
    assert zr.min() >= z.min() - 1e-6 and zr.max() <= z.max() + 1e-6


@pytest.mark.parametrize("ns,ni", [(64, 64), (128, 128), (64, 128), (128, 64)])
@pytest.mark.parametrize("mode", ["u", "det", "sorted_u"])
def test_sample_pdf_unrolled_kernel_matches_general_kernel(ns, ni, mode):
    # 64 / 128 samples run the unrolled kernel; a u that is not 16-byte aligned sends the same call to the general one
    gen = torch.Generator().manual_seed(21)
    N = 777
    z = torch.sort(torch.rand(N, ns, generator=gen) * 4 + 2, dim=-1)[0].to(DEV)
    w = torch.rand(N, ns, generator=gen) ** 6
    w[5] = 0.0
    w[6, 3:] = 0.0
    w = w.to(DEV)
    u = torch.rand(N, ni, generator=gen)
    u[7, -1] = 1.0
    u[8, 0] = 0.0
    if mode == "sorted_u":
        u = torch.sort(u, -1)[0]
    buf = torch.empty(N * ni + 1, device=DEV)
    buf[1:] = u.reshape(-1).to(DEV)
    u_odd = buf[1:].view(N, ni)
    assert u_odd.data_ptr() % 16 == 4 and u_odd.is_contiguous()
    det = mode == "det"
    fast = K.sample_pdf_merge(z, w, ni, det, u.to(DEV), debug=True)
    slow = K.sample_pdf_merge(z, w, ni, det, u_odd, debug=True)
    for a, b, name in zip(fast, slow, ("z", "z_new", "inds")):
        assert torch.equal(a, b), name
    # in-kernel Philox stream: same counters, same words
    if mode == "u":
        assert torch.equal(K.sample_pdf_merge(z, w, ni, False, None, seed=11), _general_philox(z, w, ni, 11))


def _general_philox(z, w, ni, seed):
    # the general kernel with its own draws: an output buffer that is only 4-byte aligned
    import ctypes as C
    from nerf_fl_b200 import _lib as L
    n, ns = z.shape
    buf = torch.empty(n * (ns + ni) + 1, device=z.device)
    out = buf[1:].view(n, ns + ni)
    L.check(L.lib().nwb_sample_pdf_merge(L.ptr(z), L.ptr(w), n, ns, ni, 0, None, seed, None, L.ptr(out), None, None,
                                         L.stream_ptr(z.device)), "nwb_sample_pdf_merge")
    return out


def test_sample_pdf_golden_det(golden):
    # the committed golden (reference sample_pdf) pins det-mode u == torch.linspace bit for bit
    g = golden("sample_pdf")
    z, w, _ = _pdf_inputs(None)
    _, _, ii = K.sample_pdf_merge(z.to(DEV), w.to(DEV), 64, True, None, debug=True)
    mid = 0.5 * (z[:, :-1] + z[:, 1:])
    _, cdf, uu, inds = O.sample_pdf(mid, w[:, 1:-1], 64, det=True, return_aux=True)
    assert torch.equal(uu, g["u_det"])
    # own CDF: flips only where u ties a cdf value to the last ulp -- u == 1.0 vs cdf[-1] ~ 1.0
    # (last column) does so in every row and is value-neutral (both branches give bins[-1])
    assert (ii.cpu()[:, :-1] != inds[:, :-1]).float().mean().item() < 2e-3


# ---- (a5) compositing --------------------------------------------------------------------------
@pytest.mark.parametrize("transient", [False, True])
@pytest.mark.parametrize("white_back", [False, True])
@pytest.mark.parametrize("test_time", [False, True])
def test_composite_fwd_golden(golden, transient, white_back, test_time):
    g = golden("composite")
    tag = f"t{int(transient)}_w{int(white_back)}_x{int(test_time)}/"
    typ = "fine" if transient else "coarse"
    raw = (g["raw"] if transient else g["raw"][..., :4]).contiguous()
    got = K.composite_fwd(raw.to(DEV), g["z"].to(DEV), typ=typ, transient=transient,
                          white_back=white_back, test_time=test_time, beta_min=0.1,
                          noise=g["noise"].to(DEV), noise_std=0.7)
    want = {k[len(tag):]: v for k, v in g.items() if k.startswith(tag)}
    assert set(got) == set(want)
    for k in want:
        assert maxerr(got[k], want[k]) <= 2e-6, k


def test_composite_sigma_only():
    gen = torch.Generator().manual_seed(4)
    z = torch.sort(torch.rand(77, 100, generator=gen) * 4 + 2, dim=-1)[0]
    raw = torch.rand(77, 100, 1, generator=gen) ** 3 * 20
    want = O.composite(raw, z, typ="coarse", transient=False, white_back=True, test_time=True)
    got = K.composite_fwd(raw.to(DEV), z.to(DEV), typ="coarse", transient=False, white_back=True,
                          test_time=True)
    assert set(got) == set(want) == {"weights_coarse", "opacity_coarse"}
    for k in want:
        assert maxerr(got[k], want[k]) <= 5e-6, k


@pytest.mark.parametrize("S,C", [(1, 4), (31, 9), (64, 4), (128, 9), (256, 9), (257, 4)])
def test_composite_fwd_bwd_vs_oracle(S, C):
    gen = torch.Generator().manual_seed(S * 10 + C)
    N = 203
    z = torch.sort(torch.rand(N, S, generator=gen) * 4 + 2, dim=-1)[0]
    raw = torch.rand(N, S, C, generator=gen)
    raw[..., 3 if C > 1 else 0] = raw[..., 3 if C > 1 else 0] ** 3 * 20
    if C == 9:
        raw[..., 7] = raw[..., 7] ** 3 * 6
    transient = C == 9
    typ = "fine" if transient else "coarse"
    noise = torch.randn(N, S, generator=gen)
    raw_o = raw.clone().requires_grad_(True)
    want = O.composite(raw_o, z, typ=typ, transient=transient, white_back=True, test_time=False,
                       beta_min=0.03, noise=noise, noise_std=0.5)
    got = K.composite_fwd(raw.to(DEV), z.to(DEV), typ=typ, transient=transient, white_back=True,
                          test_time=False, beta_min=0.03, noise=noise.to(DEV), noise_std=0.5)
    assert set(got) == set(want)
    for k in want:
        assert maxerr(got[k], want[k]) <= 5e-6, k
    # backward: random cotangents on every differentiable output
    cot = {k: torch.randn(v.shape, generator=gen) for k, v in want.items()}
    loss = sum((want[k] * cot[k]).sum() for k in want)
    (d_ref,) = torch.autograd.grad(loss, raw_o)
    d_got = K.composite_bwd(raw.to(DEV), z.to(DEV), {k: v.to(DEV) for k, v in cot.items()}, typ=typ,
                            transient=transient, white_back=True, noise=noise.to(DEV), noise_std=0.5)
    scale = d_ref.abs().max().item()
    assert maxerr(d_got, d_ref) <= 2e-5 * max(scale, 1.0), (maxerr(d_got, d_ref), scale)


@pytest.mark.parametrize("S,C,transient", [(1, 4, False), (33, 9, True), (64, 4, False), (100, 9, False), (128, 9, True),
                                          (128, 4, False), (200, 9, True), (256, 9, True), (256, 4, False)])
def test_composite_bwd_resident_kernel_matches_ring_kernel(S, C, transient):
    # nwb_composite_bwd_select: 2 = the whole-ray-resident backward up to S = 256, 1 = the ring kernel
    gen = torch.Generator().manual_seed(S * 7 + C)
    N = 517
    z = torch.sort(torch.rand(N, S, generator=gen) * 4 + 2, dim=-1)[0].to(DEV)
    raw = torch.rand(N, S, C, generator=gen)
    raw[..., 3] = raw[..., 3] ** 3 * 20 - (0.0 if transient else 1.0)      # static-only passes clamp sigma + noise at 0
    if C == 9:
        raw[..., 7] = raw[..., 7] ** 3 * 6
    raw = raw.to(DEV)
    typ = "fine" if C == 9 else "coarse"
    noise = None if transient else torch.randn(N, S, generator=gen).to(DEV)
    keys = [f"weights_{typ}", f"opacity_{typ}", f"depth_{typ}", "rgb_fine" if transient else f"rgb_{typ}"]
    shapes = {f"weights_{typ}": (N, S), f"opacity_{typ}": (N,), f"depth_{typ}": (N,), "rgb_fine": (N, 3), f"rgb_{typ}": (N, 3),
              "transient_sigmas": (N, S), "beta": (N,), "_rgb_fine_static": (N, 3), "_rgb_fine_transient": (N, 3)}
    if transient:
        keys += ["transient_sigmas", "beta", "_rgb_fine_static", "_rgb_fine_transient"]
    cot = {k: torch.randn(shapes[k], generator=gen).to(DEV) for k in keys}
    kw = dict(typ=typ, transient=transient, white_back=True, noise=noise, noise_std=0.5)
    amax_a, amax_b = torch.zeros(64, dtype=torch.int32, device=DEV), torch.zeros(64, dtype=torch.int32, device=DEV)
    try:
        K.composite_bwd_select(2)
        a = K.composite_bwd(raw, z, cot, amax_out=amax_a, **kw)
        K.composite_bwd_select(1)
        b = K.composite_bwd(raw, z, cot, amax_out=amax_b, **kw)
    finally:
        K.composite_bwd_select(0)
    scale = b.abs().max().item()
    assert maxerr(a, b) <= 2e-6 * max(scale, 1.0), (maxerr(a, b), scale)
    # max|d_raw| as fp32 bits, from either kernel, is the maximum of what it wrote
    assert amax_a[0].item() == a.abs().max().view(torch.int32).item()
    assert amax_b[0].item() == b.abs().max().view(torch.int32).item()


def test_composite_empty():
    out = K.composite_fwd(torch.zeros(0, 8, 4, device=DEV), torch.zeros(0, 8, device=DEV), typ="coarse",
                          transient=False, white_back=False, test_time=False)
    assert out["rgb_coarse"].shape == (0, 3)


# ---- tcgen05 tile engine -----------------------------------------------------------------------
@pytest.mark.parametrize("N,Kd", [(256, 256), (256, 64), (128, 256), (128, 128), (16, 64), (256, 192)])
def test_umma_k_major(N, Kd):
    gen = torch.Generator().manual_seed(N + Kd)
    A = torch.randn(128, Kd, generator=gen).to(torch.bfloat16)
    B = torch.randn(N, Kd, generator=gen).to(torch.bfloat16)
    d = K.selftest_umma(layout.tile_image(A).to(DEV), layout.tile_image(B).to(DEV), N, Kd, 0)
    want = A.float() @ B.float().T
    assert maxerr(d, want) <= 1e-3 * Kd ** 0.5, maxerr(d, want)


@pytest.mark.parametrize("N,Kd", [(256, 128), (128, 128), (64, 64), (256, 16)])
def test_umma_mn_major(N, Kd):
    gen = torch.Generator().manual_seed(N * 3 + Kd)
    X = torch.randn(Kd, 128, generator=gen).to(torch.bfloat16)     # contraction index = rows
    Y = torch.randn(Kd, N, generator=gen).to(torch.bfloat16)
    d = K.selftest_umma(layout.tile_image(X).to(DEV), layout.tile_image(Y).to(DEV), N, Kd, 1)
    want = X.float().T @ Y.float()
    assert maxerr(d, want) <= 1e-3 * Kd ** 0.5, maxerr(d, want)


@pytest.mark.parametrize("N,Kd", [(256, 256), (128, 128), (64, 64), (256, 128)])
def test_umma_a_kmajor_b_mnmajor(N, Kd):
    gen = torch.Generator().manual_seed(N * 5 + Kd)
    A = torch.randn(128, Kd, generator=gen).to(torch.bfloat16)
    Wt = torch.randn(Kd, N, generator=gen).to(torch.bfloat16)       # contraction index = rows
    d = K.selftest_umma(layout.tile_image(A).to(DEV), layout.tile_image(Wt).to(DEV), N, Kd, 2)
    want = A.float() @ Wt.float()
    assert maxerr(d, want) <= 1e-3 * Kd ** 0.5, maxerr(d, want)


@pytest.mark.parametrize("N,Kd", [(256, 128), (128, 128), (256, 32), (128, 64)])
def test_umma_f8_mn_major(N, Kd):
    """The 8-bit weight-gradient shape: D = A^T . B with A (K x 128) e5m2 and B (K x N) e4m3, both read MN-major from
    8-bit tile images (tcgen05.mma kind::f8f6f4, K = 32 per instruction, fp32 accumulate: products are exact)."""
    gen = torch.Generator().manual_seed(N * 7 + Kd)
    X = torch.randn(Kd, 128, generator=gen).to(torch.float8_e5m2)
    Y = torch.randn(Kd, N, generator=gen).to(torch.float8_e4m3fn)
    d = K.selftest_umma(layout.tile_image8(X.float(), torch.float8_e5m2).to(DEV),
                        layout.tile_image8(Y.float(), torch.float8_e4m3fn).to(DEV), N, Kd, 3)
    want = X.float().T @ Y.float()
    assert maxerr(d, want) <= 1e-4 * Kd ** 0.5, maxerr(d, want)


@pytest.mark.parametrize("N,Kd", [(256, 128), (128, 128), (256, 32), (96, 128), (32, 64)])
def test_umma_f8_mn_major_swizzle32(N, Kd):
    """The same contraction out of SWIZZLE_32B images -- the layout the field kernels store (32-column groups of rows x 32 B:
    a warp of 32 rows writes 1 KB of contiguous bytes from registers)."""
    gen = torch.Generator().manual_seed(N * 11 + Kd)
    X = torch.randn(Kd, 128, generator=gen).to(torch.float8_e5m2)
    Y = torch.randn(Kd, N, generator=gen).to(torch.float8_e4m3fn)
    d = K.selftest_umma(layout.tile_image8s32(X.float(), torch.float8_e5m2).to(DEV),
                        layout.tile_image8s32(Y.float(), torch.float8_e4m3fn).to(DEV), N, Kd, 4)
    want = X.float().T @ Y.float()
    assert maxerr(d, want) <= 1e-4 * Kd ** 0.5, maxerr(d, want)


# ---- fused field forward (a3, a4) --------------------------------------------------------------
def _sd_dev(sd):
    return {k: v.to(DEV).contiguous() for k, v in sd.items()}


@pytest.mark.parametrize("head,scale", [("sigma_only", 1.0), ("plain", 1.0), ("plain", 3.0), ("a", 1.0),
                                        ("at", 1.0), ("at", 3.0)])
def test_field_forward_golden(golden, head, scale):
    g = golden("field")
    x = g["x"]
    tag = f"{head}_s{int(scale)}"
    enc_a, enc_t = head in ("a", "at"), head == "at"
    seed = {"sigma_only": 21, "plain": 21, "a": 22, "at": 23}[head]
    typ = "coarse" if head in ("sigma_only", "plain") else "fine"
    sd = syn.make_field_params(seed, typ, enc_a, enc_t, weight_scale=scale)
    desc = K.field_desc(63, 27, 48 if enc_a else 0, 16 if enc_t else 0)
    sdd = _sd_dev(sd)
    packed = K.field_pack(desc, sdd)
    B = x.shape[0]
    xd = x.to(DEV)
    mode = K.mlp_mode(sigma_only=head == "sigma_only", transient=enc_t)
    rb = None
    if head != "sigma_only":
        _, rb, _ = K.ray_consts_fwd(desc, sdd, xd[:, 63:90].contiguous(), dir_is_encoded=True,
                                    a_rows=xd[:, 90:138].contiguous() if enc_a else None,
                                    t_rows=xd[:, 138:154].contiguous() if enc_t else None, transient=enc_t)
    got = K.mlp_fwd(desc, packed, n_rays=B, S=1, mode=mode, ray_bias=rb, x_enc=xd[:, :63].contiguous())
    want = g[tag]
    assert got.shape == want.shape
    # bf16 operands, fp32 accumulate: tolerance of BASELINE.json (1e-3 on composited outputs) scaled for raw heads
    tol = 3e-3 if scale == 1.0 else 3e-2
    assert maxerr(got, want) <= tol * max(1.0, want.abs().max().item()), maxerr(got, want)


@pytest.mark.parametrize("n_rays,S,transient", [(5, 64, False), (37, 128, True), (300, 64, True), (1, 7, True)])
def test_field_forward_rays_vs_oracle(n_rays, S, transient):
    gen = torch.Generator().manual_seed(n_rays + S)
    rays = syn.blender_rays(n_rays, seed=3)
    z = torch.sort(torch.rand(n_rays, S, generator=gen) * 4 + 2, dim=-1)[0]
    sd = syn.make_field_params(31, "fine", True, transient)
    emb_a, emb_t = syn.make_embedding(5, 20, 48), syn.make_embedding(6, 20, 16)
    ts = torch.randint(0, 20, (n_rays,), generator=gen)
    desc = K.field_desc(63, 27, 48, 16 if transient else 0)
    sdd = _sd_dev(sd)
    packed = K.field_pack(desc, sdd)
    _, rb, err = K.ray_consts_fwd(desc, sdd, rays[:, 3:6].to(DEV).contiguous(), n_freq_dir=4, ts=ts.to(DEV),
                                  emb_a=emb_a.to(DEV), emb_t=emb_t.to(DEV) if transient else None,
                                  transient=transient)
    assert err.item() == 0
    mode = K.mlp_mode(transient=transient, save=True)
    saved = K.aligned_empty(K.mlp_save_bytes(desc, n_rays * S, mode), DEV, 16)
    got = K.mlp_fwd(desc, packed, n_rays=n_rays, S=S, mode=mode, ray_bias=rb, rays=rays.to(DEV), z=z.to(DEV),
                    n_freq_xyz=10, saved=saved)
    got2 = K.mlp_fwd(desc, packed, n_rays=n_rays, S=S, mode=K.mlp_mode(transient=transient), ray_bias=rb,
                     rays=rays.to(DEV), z=z.to(DEV), n_freq_xyz=10)
    assert torch.equal(got, got2)                      # saving activations does not change the result
    xyz = (rays[:, None, 0:3] + rays[:, None, 3:6] * z[:, :, None]).reshape(-1, 3)
    parts = [O.posenc(xyz, 10), O.posenc(rays[:, 3:6], 4).repeat_interleave(S, 0), emb_a[ts].repeat_interleave(S, 0)]
    if transient:
        parts.append(emb_t[ts].repeat_interleave(S, 0))
    want = O.field_forward(sd, torch.cat(parts, -1), in_dir_a=27 + 48, output_transient=transient)
    assert maxerr(got, want) <= 3e-3 * max(1.0, want.abs().max().item()), maxerr(got, want)


def _relerr(got, ref):
    got, ref = got.detach().double().cpu().flatten(), ref.detach().double().cpu().flatten()
    l2 = (got - ref).norm().item() / max(ref.norm().item(), 1e-30)
    cos = torch.nn.functional.cosine_similarity(got, ref, dim=0).item()
    return l2, cos


@pytest.mark.parametrize("n_rays,S,transient,scale", [(6, 64, True, 1.0), (41, 128, True, 3.0), (9, 64, False, 1.0),
                                                     (130, 96, True, 1.0)])
def test_field_backward_vs_oracle(n_rays, S, transient, scale):
    """Every parameter / embedding gradient of one field evaluation, with random cotangents, against
    autograd through (i) the oracle's operand-rounding model of the kernels (bf16 GEMM operands, 8-bit images in the
    weight-gradient contraction: O.field_forward_q8) and (ii) the plain fp32 oracle.
    Per-sample gradients of a ReLU net are discontinuous in the forward's rounding (a gate whose
    pre-activation sits within 2^-9 of zero opens in one forward and not the other), so with bf16
    operands the fp32 comparison is bounded by gate noise, not by the kernel: (i) is the tight check,
    (ii) bounds direction and size.  Head gradients (no gate downstream) must meet 1e-2 in both."""
    gen = torch.Generator().manual_seed(n_rays * 7 + S)
    rays = syn.blender_rays(n_rays, seed=3)
    z = torch.sort(torch.rand(n_rays, S, generator=gen) * 4 + 2, dim=-1)[0]
    sd = syn.make_field_params(33, "fine", True, transient, weight_scale=scale)
    n_vocab = 7
    emb_a, emb_t = syn.make_embedding(5, n_vocab, 48), syn.make_embedding(6, n_vocab, 16)
    ts = torch.randint(0, n_vocab, (n_rays,), generator=gen)
    C9 = 9 if transient else 4
    d_raw = torch.randn(n_rays * S, C9, generator=gen)

    def oracle_grads(emulate):
        sd_o = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
        ea_o, et_o = emb_a.clone().requires_grad_(True), emb_t.clone().requires_grad_(True)
        xyz = (rays[:, None, 0:3] + rays[:, None, 3:6] * z[:, :, None]).reshape(-1, 3)
        x_e = O.posenc(xyz, 10)
        dir_a = torch.cat([O.posenc(rays[:, 3:6], 4).repeat_interleave(S, 0), ea_o[ts].repeat_interleave(S, 0)], -1)
        tt = et_o[ts].repeat_interleave(S, 0) if transient else None
        if emulate:
            raw_o = O.field_forward_q8(sd_o, x_e, dir_a, tt, transient)
        else:
            raw_o = O.field_forward(sd_o, torch.cat([x_e, dir_a] + ([tt] if transient else []), -1), in_dir_a=27 + 48,
                                    output_transient=transient)
        (raw_o * d_raw).sum().backward()
        g = {k: v.grad for k, v in sd_o.items()}
        g["emb_a"] = ea_o.grad
        if transient:
            g["emb_t"] = et_o.grad
        return g

    # ---- product ----
    desc = K.field_desc(63, 27, 48, 16 if transient else 0)
    sdd = _sd_dev(sd)
    packed = K.field_pack(desc, sdd)
    ts_d, ea_d, et_d = ts.to(DEV), emb_a.to(DEV), emb_t.to(DEV)
    feat, rb, _ = K.ray_consts_fwd(desc, sdd, rays[:, 3:6].to(DEV).contiguous(), n_freq_dir=4, ts=ts_d, emb_a=ea_d,
                                   emb_t=et_d if transient else None, transient=transient)
    mode = K.mlp_mode(transient=transient, save=True)
    saved = K.aligned_empty(K.mlp_save_bytes(desc, n_rays * S, mode), DEV, 16)
    raw = K.mlp_fwd(desc, packed, n_rays=n_rays, S=S, mode=mode, ray_bias=rb, rays=rays.to(DEV), z=z.to(DEV), saved=saved)
    grads = {k: torch.zeros_like(v) for k, v in sdd.items()}
    d_rb = torch.zeros(n_rays, 256, device=DEV)
    K.mlp_bwd(desc, packed, raw, d_raw.to(DEV), saved, n_rays=n_rays, S=S, mode=mode, grads=grads, d_ray_bias=d_rb)
    d_ea, d_et = torch.zeros_like(ea_d), torch.zeros_like(et_d)
    K.ray_consts_bwd(desc, sdd, feat, d_rb, ts=ts_d, n_vocab=n_vocab, transient=transient, grads=grads, d_emb_a=d_ea,
                     d_emb_t=d_et if transient else None)
    torch.cuda.synchronize()
    grads["emb_a"] = d_ea
    if transient:
        grads["emb_t"] = d_et
    g_emu, g_f32 = oracle_grads(True), oracle_grads(False)
    bad = []
    for k in g_f32:
        l2e, cose = _relerr(grads[k], g_emu[k])
        l2f, cosf = _relerr(grads[k], g_f32[k])
        ungated = k.startswith(("static_sigma", "static_rgb", "transient_sigma", "transient_rgb", "transient_beta")) \
            and k.endswith("bias")
        if ungated:
            # scalar sums of signed terms: ill-conditioned under the x3 weights' larger forward differences
            ok = (l2e <= 1e-2 and l2f <= 1e-2) if scale == 1.0 else l2e <= 5e-2
        else:
            ok = l2e <= 8e-2 and cose >= 0.997 and l2f <= 0.25 and cosf >= 0.97
        if not ok:
            bad.append((k, round(l2e, 4), round(cose, 5), round(l2f, 4), round(cosf, 5)))
    assert not bad, bad


def test_copy_segments_one_launch():
    g = torch.Generator().manual_seed(2)
    src = [torch.rand(1024, 8, generator=g).to(DEV), torch.randint(0, 1500, (1024,), generator=g).to(DEV),
           torch.rand(1024, 3, generator=g).to(DEV), torch.rand(7, generator=g).to(DEV),
           torch.randint(0, 255, (13,), generator=g).to(torch.uint8).to(DEV)]
    src.append(torch.rand(33, generator=g).to(DEV)[1:])                # 4-byte aligned only
    dst = [torch.zeros_like(s) for s in src]
    K.copy_segments(list(zip(dst, src)))
    for d, s in zip(dst, src):
        assert torch.equal(d, s)
    with pytest.raises(RuntimeError):
        K.copy_segments([(torch.zeros(4, device=DEV), torch.zeros(5, device=DEV))])
