"""Thin torch-tensor wrappers over the C ABI (one function per entry point).

PyTorch is used for device memory and streams only; every computation below happens inside
libnerfw_b200.so on the current CUDA stream.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib as L


# ---- optional per-call device timing (bench.py's roofline leg) -------------------------------------
_timing = None     # None, or a list collecting (name, units, mode, start_event, end_event)
launch_count = 0   # kernels the library has launched in this process (nwb_kernel_launches; bench.py's gpu_launches)


class _timed:
    """Brackets one C-ABI call with CUDA events on the launching stream when timing is on."""

    def __init__(self, name, units, mode=0, dev=None):
        self.name, self.units, self.mode = name, units, mode
        self.guard = None
        if dev is not None:
            dev = torch.device(dev)
            if dev.type == "cuda" and dev.index is not None and dev.index != torch.cuda.current_device():
                self.guard = torch.cuda.device(dev)      # kernels launch on the CURRENT device: make it the tensors'

    def __enter__(self):
        if self.guard is not None:
            self.guard.__enter__()
        if _timing is not None:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e0.record()
        return self

    def __exit__(self, *exc):
        global launch_count
        launch_count = int(_launches())
        if _timing is not None:
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            _timing.append((self.name, self.units, self.mode, self.e0, e1))
        if self.guard is not None:
            self.guard.__exit__(*exc)
        return False


_launches_fn = None


def _launches():
    global _launches_fn
    if _launches_fn is None:
        _launches_fn = L.lib().nwb_kernel_launches
        _launches_fn.restype = C.c_longlong
        _launches_fn.argtypes = []
    return _launches_fn()


def timing_start():
    global _timing
    _timing = []


def timing_stop():
    """-> {(name, mode): (launches, total_ms, total_units)}; synchronises."""
    global _timing
    rec, _timing = _timing, None
    torch.cuda.synchronize()
    out = {}
    for name, units, mode, e0, e1 in rec or []:
        n, ms, u = out.get((name, mode), (0, 0.0, 0))
        out[(name, mode)] = (n + 1, ms + e0.elapsed_time(e1), u + units)
    return out


def _f32(t: torch.Tensor) -> torch.Tensor:
    return t if (t.dtype == torch.float32 and t.is_contiguous()) else t.float().contiguous()


def sample_coarse(rays: torch.Tensor, n_samples: int, use_disp: bool = False, perturb: float = 0.0,
                  perturb_rand: torch.Tensor | None = None, seed: int = 0) -> torch.Tensor:
    """z_vals (N, n_samples) of models/rendering.py:244-259."""
    rays = _f32(rays)
    n = rays.shape[0]
    z = torch.empty(n, n_samples, device=rays.device, dtype=torch.float32)
    pr = _f32(perturb_rand) if perturb_rand is not None else None
    with _timed("nwb_sample_coarse", n, dev=rays.device):
        L.check(L.lib().nwb_sample_coarse(L.ptr(rays), rays.shape[1], n, n_samples, int(bool(use_disp)),
                                          float(perturb), L.ptr(pr), int(seed) & (2 ** 64 - 1), L.ptr(z),
                                          L.stream_ptr(rays.device)), "nwb_sample_coarse")
    return z


def posenc(x: torch.Tensor, n_freqs: int) -> torch.Tensor:
    """PosEmbedding.forward, models/nerf.py:19-32."""
    x = _f32(x)
    assert x.dim() == 2 and x.shape[1] == 3, "PosEmbedding expects (B, 3)"
    out = torch.empty(x.shape[0], 3 + 6 * n_freqs, device=x.device, dtype=torch.float32)
    with _timed("nwb_posenc", x.shape[0], dev=x.device):
        L.check(L.lib().nwb_posenc(L.ptr(x), x.shape[0], n_freqs, L.ptr(out), L.stream_ptr(x.device)),
                "nwb_posenc")
    return out


def sample_pdf_merge(z_coarse: torch.Tensor, w_coarse: torch.Tensor, n_importance: int, det: bool,
                     u: torch.Tensor | None = None, seed: int = 0, cdf: torch.Tensor | None = None,
                     debug: bool = False):
    """sort(cat[z, sample_pdf(mid(z), w[:,1:-1])]) of models/rendering.py:267-272 (+ :7-46)."""
    z_coarse, w_coarse = _f32(z_coarse), _f32(w_coarse)
    n, ns = z_coarse.shape
    dev = z_coarse.device
    z_out = torch.empty(n, ns + n_importance, device=dev, dtype=torch.float32)
    z_new = torch.empty(n, n_importance, device=dev, dtype=torch.float32) if debug else None
    inds = torch.empty(n, n_importance, device=dev, dtype=torch.int64) if debug else None
    u_ = _f32(u) if u is not None else None
    cdf_ = _f32(cdf) if cdf is not None else None
    with _timed("nwb_sample_pdf_merge", n, dev=dev):
        L.check(L.lib().nwb_sample_pdf_merge(L.ptr(z_coarse), L.ptr(w_coarse), n, ns, n_importance,
                                             int(bool(det)), L.ptr(u_), int(seed) & (2 ** 64 - 1),
                                             L.ptr(cdf_), L.ptr(z_out), L.ptr(z_new), L.ptr(inds),
                                             L.stream_ptr(dev)), "nwb_sample_pdf_merge")
    return (z_out, z_new, inds) if debug else z_out


def copy_segments(pairs) -> None:
    """[(dst, src), ...] contiguous CUDA tensors of equal byte size, <= 8 of them, copied by one launch."""
    n = len(pairs)
    srcs, dsts, nbytes = (C.c_void_p * n)(), (C.c_void_p * n)(), (C.c_int64 * n)()
    for i, (d, s_) in enumerate(pairs):
        if d.numel() * d.element_size() != s_.numel() * s_.element_size() or d.dtype != s_.dtype:
            raise RuntimeError("copy_segments: source and destination must have the same dtype and size")
        srcs[i], dsts[i], nbytes[i] = L.ptr(s_), L.ptr(d), d.numel() * d.element_size()
    fn = _fn("nwb_copy_segments", [_i, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_int64), _vp])
    with _timed("nwb_copy_segments", n, dev=pairs[0][0].device):
        L.check(fn(n, srcs, dsts, nbytes, L.stream_ptr(pairs[0][0].device)), "nwb_copy_segments")


def composite_bwd_select(gen: int) -> None:
    """0 = whole-ray-resident compositing backward for S <= 128 (default), 1 = always the chunk-ring kernel, 2 = resident up to S = 256."""
    L.check(_fn("nwb_composite_bwd_select", [_i])(int(gen)), "nwb_composite_bwd_select")


def composite_flags(transient: bool, white_back: bool, test_time: bool, coarse: bool) -> int:
    return ((L.NWB_COMP_TRANSIENT if transient else 0) | (L.NWB_COMP_WHITE_BACK if white_back else 0)
            | (L.NWB_COMP_TEST_TIME if test_time else 0) | (L.NWB_COMP_COARSE if coarse else 0))


def composite_fwd(raw: torch.Tensor, z: torch.Tensor, *, typ: str, transient: bool, white_back: bool,
                  test_time: bool, beta_min: float = 0.0, noise: torch.Tensor | None = None,
                  noise_std: float = 0.0, clear_word: torch.Tensor | None = None) -> dict:
    """One inference() compositing pass (models/rendering.py:141-226) -> result-dict entries."""
    raw, z = _f32(raw), _f32(z)
    n, S = z.shape
    Cc = raw.shape[-1]
    dev = z.device
    new = lambda *s: torch.empty(*s, device=dev, dtype=torch.float32)  # noqa: E731
    coarse = typ == "coarse"
    res, o = {}, L.CompositeOut()
    res[f"weights_{typ}"] = new(n, S)
    res[f"opacity_{typ}"] = new(n)
    o.weights, o.opacity = L.ptr(res[f"weights_{typ}"]), L.ptr(res[f"opacity_{typ}"])
    if transient:
        res["transient_sigmas"] = new(n, S)
        o.transient_sigmas = L.ptr(res["transient_sigmas"])
    if not (test_time and coarse) and Cc != 1:
        if transient:
            res["beta"] = new(n)
            res["_rgb_fine_static"], res["_rgb_fine_transient"] = new(n, 3), new(n, 3)
            res["rgb_fine"] = new(n, 3)
            o.beta, o.rgb = L.ptr(res["beta"]), L.ptr(res["rgb_fine"])
            o.rgb_static, o.rgb_transient = L.ptr(res["_rgb_fine_static"]), L.ptr(res["_rgb_fine_transient"])
            if test_time:
                for k, shape, slot in (("rgb_fine_static", (n, 3), "tt_rgb_static"),
                                       ("depth_fine_static", (n,), "tt_depth_static"),
                                       ("rgb_fine_transient", (n, 3), "tt_rgb_transient"),
                                       ("depth_fine_transient", (n,), "tt_depth_transient")):
                    res[k] = new(*shape)
                    setattr(o, slot, L.ptr(res[k]))
        else:
            res[f"rgb_{typ}"] = new(n, 3)
            o.rgb = L.ptr(res[f"rgb_{typ}"])
        res[f"depth_{typ}"] = new(n)
        o.depth = L.ptr(res[f"depth_{typ}"])
    nz = _f32(noise) if (noise is not None and not transient and noise_std != 0.0) else None
    if clear_word is not None:
        o.clear_word = L.ptr(clear_word)         # first word <- 0 (the backward workspace's amax word, see composite_bwd)
    with _timed("nwb_composite_fwd", n, Cc * 1000 + S, dev=dev):
        L.check(L.lib().nwb_composite_fwd(L.ptr(raw), Cc, L.ptr(z), L.ptr(nz), float(noise_std), n, S,
                                          composite_flags(transient, white_back, test_time, coarse),
                                          float(beta_min), C.byref(o), L.stream_ptr(dev)),
                "nwb_composite_fwd")
    return res


def composite_bwd(raw: torch.Tensor, z: torch.Tensor, grads: dict, *, typ: str, transient: bool,
                  white_back: bool, noise: torch.Tensor | None = None,
                  noise_std: float = 0.0, seed: dict | None = None, amax_out: torch.Tensor | None = None,
                  amax_cleared: bool = False) -> torch.Tensor:
    """Cotangents keyed like the result dict -> d_raw (N,S,C).

    ``seed`` (nwb_composite_bwd_seeded): dict(rgb, beta, targets, g_terms (5 device scalars or None), coef, lambda_u,
    n_rays, n_ts, fine) -- the NerfWLoss cotangents of this pass's rgb / beta / transient_sigmas are formed in the kernel
    and added to ``grads``.  ``amax_out``: buffer whose first word receives max|d_raw| (fp32 bits) for
    mlp_bwd(mode | NWB_MLP_DZ_AMAX_GIVEN)."""
    raw, z = _f32(raw), _f32(z)
    n, S = z.shape
    Cc = raw.shape[-1]
    g = L.CompositeGrads()
    keep = []

    def put(slot, key):
        t = grads.get(key)
        if t is not None:
            t = _f32(t)
            keep.append(t)
            setattr(g, slot, L.ptr(t))

    put("d_weights", f"weights_{typ}")
    put("d_opacity", f"opacity_{typ}")
    put("d_depth", f"depth_{typ}")
    if transient:
        put("d_rgb", "rgb_fine")
        put("d_transient_sigmas", "transient_sigmas")
        put("d_beta", "beta")
        put("d_rgb_static", "_rgb_fine_static")
        put("d_rgb_transient", "_rgb_fine_transient")
    else:
        put("d_rgb", f"rgb_{typ}")
    d_raw = torch.empty_like(raw)
    nz = _f32(noise) if (noise is not None and not transient and noise_std != 0.0) else None
    flags = composite_flags(transient, white_back, False, typ == "coarse")
    if amax_cleared:
        flags |= L.NWB_COMP_AMAX_CLEARED          # composite_fwd(clear_word=amax_out) of the same pass already zeroed it
    if seed is None and amax_out is None:
        with _timed("nwb_composite_bwd", n, Cc * 1000 + S, dev=z.device):
            L.check(L.lib().nwb_composite_bwd(L.ptr(raw), Cc, L.ptr(z), L.ptr(nz), float(noise_std), n, S, flags,
                                              C.byref(g), L.ptr(d_raw), L.stream_ptr(z.device)),
                    "nwb_composite_bwd")
        return d_raw
    sd = None
    if seed is not None:
        sd = L.LossSeed()
        for k in ("rgb", "beta", "targets"):
            t = seed.get(k)
            if t is not None:
                t = _f32(t)
                keep.append(t)
                setattr(sd, k, L.ptr(t))
        for i, t in enumerate(seed.get("g_terms") or ()):
            if t is not None:
                t = _f32(t)
                keep.append(t)
                sd.g_terms[i] = L.ptr(t)
        sd.coef, sd.lambda_u = float(seed["coef"]), float(seed["lambda_u"])
        sd.n_rays, sd.n_ts, sd.fine = int(seed["n_rays"]), int(seed.get("n_ts") or 0), int(bool(seed["fine"]))
    with _timed("nwb_composite_bwd", n, Cc * 1000 + S, dev=z.device):
        L.check(L.lib().nwb_composite_bwd_seeded(L.ptr(raw), Cc, L.ptr(z), L.ptr(nz), float(noise_std), n, S, flags,
                                                 C.byref(g), C.byref(sd) if sd is not None else None, L.ptr(amax_out),
                                                 L.ptr(d_raw), L.stream_ptr(z.device)),
                "nwb_composite_bwd_seeded")
    return d_raw


def selftest_umma(a_img: torch.Tensor, b_img: torch.Tensor, N: int, K: int, variant: int,
                  overrides=None) -> torch.Tensor:
    d = torch.empty(128, N, device=a_img.device, dtype=torch.float32)
    ov = None
    if overrides is not None:
        ov = (C.c_uint32 * 11)(*[int(v) for v in overrides])
    L.check(L.lib().nwb_selftest_umma(L.ptr(a_img), L.ptr(b_img), N, K, variant, ov, L.ptr(d),
                                      L.stream_ptr(a_img.device)), "nwb_selftest_umma")
    return d


# ---- field network (a3, a4) ----------------------------------------------------------------------
_i, _u32, _i64, _sz = C.c_int, C.c_uint32, C.c_int64, C.c_size_t
_vp = C.c_void_p
_bound = {}


def _fn(name, argtypes, restype=C.c_int):
    if name not in _bound:
        _bound[name] = L.bind(name, argtypes, restype)
    return _bound[name]


def aligned_empty(nbytes: int, device, align: int = 1024) -> torch.Tensor:
    """uint8 CUDA buffer whose data_ptr is `align`-byte aligned (tile images need 1 KB)."""
    raw = torch.empty(nbytes + align, dtype=torch.uint8, device=device)
    off = (-raw.data_ptr()) % align
    return raw[off:off + nbytes]


def field_desc(in_xyz=63, in_dir=27, in_a=0, in_t=0) -> L.FieldDesc:
    return L.FieldDesc(in_xyz, in_dir, in_a, in_t)


_SLOT_KEYS = ([(f"trunk_w", i, f"xyz_encoding_{i + 1}.0.weight") for i in range(8)] +
              [(f"trunk_b", i, f"xyz_encoding_{i + 1}.0.bias") for i in range(8)] +
              [("final_w", None, "xyz_encoding_final.weight"), ("final_b", None, "xyz_encoding_final.bias"),
               ("dir_w", None, "dir_encoding.0.weight"), ("dir_b", None, "dir_encoding.0.bias"),
               ("sigma_w", None, "static_sigma.0.weight"), ("sigma_b", None, "static_sigma.0.bias"),
               ("rgb_w", None, "static_rgb.0.weight"), ("rgb_b", None, "static_rgb.0.bias")] +
              [("tenc_w", j, f"transient_encoding.{2 * j}.weight") for j in range(4)] +
              [("tenc_b", j, f"transient_encoding.{2 * j}.bias") for j in range(4)] +
              [("tsigma_w", None, "transient_sigma.0.weight"), ("tsigma_b", None, "transient_sigma.0.bias"),
               ("trgb_w", None, "transient_rgb.0.weight"), ("trgb_b", None, "transient_rgb.0.bias"),
               ("tbeta_w", None, "transient_beta.0.weight"), ("tbeta_b", None, "transient_beta.0.bias")])


def field_slots(sd: dict, cls=L.FieldParams):
    """state_dict-keyed tensors (reference names) -> the C struct of device pointers.
    Returns (struct, keepalive list)."""
    st, keep = cls(), []
    for slot, idx, key in _SLOT_KEYS:
        t = sd.get(key)
        if t is None:
            continue
        assert t.dtype == torch.float32 and t.is_cuda and t.is_contiguous(), key
        keep.append(t)
        if idx is None:
            setattr(st, slot, t.data_ptr())
        else:
            getattr(st, slot)[idx] = t.data_ptr()
    return st, keep


def field_pack(desc: L.FieldDesc, sd: dict, out: torch.Tensor | None = None) -> torch.Tensor:
    """fp32 nn.Linear parameters -> packed operand cache (bf16 tile images + fp32 constants)."""
    dev = sd["xyz_encoding_1.0.weight"].device
    nbytes = _fn("nwb_field_pack_bytes", [C.POINTER(L.FieldDesc)], _sz)(C.byref(desc))
    if nbytes == 0:
        L.check(-1, "nwb_field_pack_bytes")
    if out is None:
        out = aligned_empty(nbytes, dev)
    st, keep = field_slots(sd)
    with _timed("nwb_field_pack", 1, dev=dev):
        L.check(_fn("nwb_field_pack", [C.POINTER(L.FieldDesc), C.POINTER(L.FieldParams), _vp, _vp])(
            C.byref(desc), C.byref(st), L.ptr(out), L.stream_ptr(dev)), "nwb_field_pack")
    return out


def ray_consts_fwd(desc, sd, dirs, *, dir_is_encoded=False, n_freq_dir=4, ts=None, emb_a=None, emb_t=None,
                   a_rows=None, t_rows=None, transient=False, err_flag=None):
    """Per-ray [dir|a|t] feature rows and the per-ray bias of dir_encoding / transient_encoding.0."""
    dirs = _f32(dirs)
    n = dirs.shape[0]
    dev = dirs.device
    mode = L.NWB_MLP_TRANSIENT if transient else 0
    feat = torch.empty(n, desc.in_dir + desc.in_a + desc.in_t, device=dev, dtype=torch.float32)
    bias = torch.empty(n, 256, device=dev, dtype=torch.float32)
    if err_flag is None:
        err_flag = torch.zeros(1, dtype=torch.int32, device=dev)
    st, keep = field_slots(sd)
    n_vocab = emb_a.shape[0] if emb_a is not None else (emb_t.shape[0] if emb_t is not None else 0)
    fn = _fn("nwb_ray_consts_fwd", [C.POINTER(L.FieldDesc), C.POINTER(L.FieldParams), _vp, _i, _i, _i, _vp, _vp, _vp,
                                    _i, _vp, _vp, _i, _u32, _vp, _vp, _vp, _vp])
    with _timed("nwb_ray_consts_fwd", n, mode, dev=dev):
        L.check(fn(C.byref(desc), C.byref(st), L.ptr(dirs), dirs.shape[1], int(dir_is_encoded), n_freq_dir, L.ptr(ts),
                   L.ptr(emb_a), L.ptr(emb_t), n_vocab, L.ptr(a_rows), L.ptr(t_rows), n, mode, L.ptr(feat), L.ptr(bias),
                   L.ptr(err_flag), L.stream_ptr(dev)), "nwb_ray_consts_fwd")
    return feat, bias, err_flag


def mlp_mode(sigma_only=False, transient=False, save=False) -> int:
    return ((L.NWB_MLP_SIGMA_ONLY if sigma_only else 0) | (L.NWB_MLP_TRANSIENT if transient else 0)
            | (L.NWB_MLP_SAVE if save else 0))


def mlp_save_bytes(desc, n_points: int, mode: int) -> int:
    return _fn("nwb_mlp_save_bytes", [C.POINTER(L.FieldDesc), _i64, _u32], _sz)(C.byref(desc), n_points, mode)


def mlp_fwd(desc, packed, *, n_rays, S, mode, ray_bias=None, rays=None, z=None, x_enc=None, n_freq_xyz=10,
            saved=None, out=None):
    """Fused field forward over n_rays*S points -> raw (n_points, 1|4|9)."""
    dev = packed.device
    Cc = 1 if (mode & L.NWB_MLP_SIGMA_ONLY) else (9 if (mode & L.NWB_MLP_TRANSIENT) else 4)
    if out is None:
        out = torch.empty(n_rays * S, Cc, device=dev, dtype=torch.float32)
    fn = _fn("nwb_mlp_fwd", [C.POINTER(L.FieldDesc), _vp, _vp, _i, _vp, _vp, _i, _vp, _i, _i, _u32, _vp, _vp, _vp])
    with _timed("nwb_mlp_fwd", n_rays * S, mode, dev=dev):
        L.check(fn(C.byref(desc), L.ptr(packed), L.ptr(rays), rays.shape[1] if rays is not None else 0, L.ptr(z),
                   L.ptr(x_enc), n_freq_xyz, L.ptr(ray_bias), n_rays, S, mode, L.ptr(out), L.ptr(saved),
                   L.stream_ptr(dev)), "nwb_mlp_fwd")
    return out


def field_traffic_bytes_per_point(kind: str, fine: bool) -> float:
    """HBM bytes per sample point one field kernel launch moves by design, from the library's own size queries (the
    per-tile record sizes of nwb_mlp_save_bytes / nwb_mlp_bwd_workspace_bytes): fwd (saving) writes the saved record;
    dgrad writes the dZ record and reads the sign masks; wgrad reads both records (without the masks) once, and the
    dpre block once per head unit."""
    desc = field_desc(63, 27, 48 if fine else 0, 16 if fine else 0)
    mode = (L.NWB_MLP_TRANSIENT if fine else 0) | L.NWB_MLP_SAVE
    n = 128 * 1024                                   # per-point slope: the workspace's constant parts (header, wgrad partial slots) drop out
    save = (mlp_save_bytes(desc, 2 * n, mode) - mlp_save_bytes(desc, n, mode)) / float(n)
    ws = (mlp_bwd_workspace_bytes(desc, 2 * n, mode & ~L.NWB_MLP_SAVE) - mlp_bwd_workspace_bytes(desc, n, mode & ~L.NWB_MLP_SAVE)) / float(n)
    masks = 13 * 4096 / 128.0
    dpre_again = ((4 if fine else 3) - 1) * 128.0     # every head unit of wgrad reads the dpre block (16 KB per tile)
    return {"fwd": save, "dgrad": ws + masks, "wgrad": (save - masks) + ws + dpre_again}[kind]


def mlp_fwd_select(gen: int) -> None:
    """Which forward kernel nwb_mlp_fwd launches: 0 = per launch, whichever is faster (default); 1 = the one-tile
    kernel; 2 = the CTA-pair kernel (parity tests walk every dispatch branch with it)."""
    L.check(_fn("nwb_mlp_fwd_select", [_i])(int(gen)), "nwb_mlp_fwd_select")


def mlp_dgrad_select(gen: int) -> None:
    """Which dgrad kernel the backward launches: 0 / 2 = two-slot pair kernel (default); 1 = first-generation kernel (parity
    tests walk both)."""
    L.check(_fn("nwb_mlp_dgrad_select", [_i])(int(gen)), "nwb_mlp_dgrad_select")


def mlp_bwd_workspace_bytes(desc, n_points: int, mode: int) -> int:
    return _fn("nwb_mlp_bwd_workspace_bytes", [C.POINTER(L.FieldDesc), _i64, _u32], _sz)(C.byref(desc), n_points, mode)


def mlp_bwd(desc, packed, raw, d_raw, saved, *, n_rays, S, mode, grads: dict, d_ray_bias, workspace=None,
            phases=("dgrad", "wgrad")):
    """Backward of mlp_fwd: accumulates into `grads` (state_dict-keyed fp32 tensors, caller zeroes)
    and into d_ray_bias (n_rays, 256).  Two kernels (dgrad chain, then wgrad), launched as separate
    C-ABI calls so the timing hook sees them apart."""
    dev = packed.device
    mode &= ~L.NWB_MLP_SAVE
    if workspace is None:
        workspace = aligned_empty(mlp_bwd_workspace_bytes(desc, n_rays * S, mode), dev, 16)
    st, keep = field_slots(grads, L.FieldGrads)
    sig = [C.POINTER(L.FieldDesc), _vp, _vp, _vp, _vp, _i, _i, _u32, C.POINTER(L.FieldGrads), _vp, _vp, _vp]
    for ph in phases:
        name = "nwb_mlp_" + ph
        with _timed(name, n_rays * S, mode & ~L.NWB_MLP_DZ_AMAX_GIVEN, dev=dev):
            L.check(_fn(name, sig)(C.byref(desc), L.ptr(packed), L.ptr(raw), L.ptr(d_raw), L.ptr(saved), n_rays, S, mode,
                                   C.byref(st), L.ptr(d_ray_bias), L.ptr(workspace), L.stream_ptr(dev)), name)
    return workspace


def ray_consts_bwd(desc, sd, ray_feat, d_ray_bias, *, ts=None, n_vocab=0, transient=False, grads: dict,
                   d_emb_a=None, d_emb_t=None, d_a_rows=None, d_t_rows=None):
    """d_ray_bias -> dir_encoding / transient_encoding.0 columns [256:], their biases, embedding rows."""
    dev = ray_feat.device
    mode = L.NWB_MLP_TRANSIENT if transient else 0
    st, keep = field_slots(sd)
    gt, keep2 = field_slots(grads, L.FieldGrads)
    fn = _fn("nwb_ray_consts_bwd", [C.POINTER(L.FieldDesc), C.POINTER(L.FieldParams), _vp, _vp, _vp, _i, _i, _u32,
                                    C.POINTER(L.FieldGrads), _vp, _vp, _vp, _vp, _vp])
    with _timed("nwb_ray_consts_bwd", ray_feat.shape[0], mode, dev=dev):
        L.check(fn(C.byref(desc), C.byref(st), L.ptr(ray_feat), L.ptr(d_ray_bias), L.ptr(ts), n_vocab, ray_feat.shape[0],
                   mode, C.byref(gt), L.ptr(d_emb_a), L.ptr(d_emb_t), L.ptr(d_a_rows), L.ptr(d_t_rows),
                   L.stream_ptr(dev)), "nwb_ray_consts_bwd")


# ---- objective (f1): NerfWLoss forward / backward -------------------------------------------------
_loss_scratch = {}


def _loss_scratch_for(dev) -> torch.Tensor:
    key = (dev.type, dev.index)
    if key not in _loss_scratch:
        n = _fn("nwb_nerfw_loss_scratch_bytes", [], _sz)()
        _loss_scratch[key] = torch.zeros(n, dtype=torch.uint8, device=dev)     # zeroed once; the kernel resets its ticket
    return _loss_scratch[key]


def nerfw_loss_fwd(rgb_coarse, rgb_fine, beta, transient_sigmas, targets, lambda_u: float, coef: float,
                   out: torch.Tensor | None = None) -> torch.Tensor:
    """-> terms (5,) = [c_l, f_l, b_l, s_l, total] on the device (one launch)."""
    dev = rgb_coarse.device
    terms = out if out is not None else torch.empty(5, device=dev, dtype=torch.float32)
    n_ts = transient_sigmas.numel() if transient_sigmas is not None else 0
    fn = _fn("nwb_nerfw_loss_fwd", [_vp] * 5 + [_i, _i64, C.c_float, C.c_float, _vp, _vp, _vp])
    with _timed("nwb_nerfw_loss_fwd", rgb_coarse.shape[0], dev=dev):
        L.check(fn(L.ptr(rgb_coarse), L.ptr(rgb_fine), L.ptr(beta), L.ptr(transient_sigmas), L.ptr(targets),
                   rgb_coarse.shape[0], n_ts, lambda_u, coef, L.ptr(_loss_scratch_for(dev)), L.ptr(terms),
                   L.stream_ptr(dev)), "nwb_nerfw_loss_fwd")
    return terms


def nerfw_loss_bwd(rgb_coarse, rgb_fine, beta, transient_sigmas, targets, lambda_u: float, coef: float, g_terms):
    """g_terms: 5 device scalars or None -> (d_rgb_coarse, d_rgb_fine, d_beta, d_transient_sigmas) (one launch)."""
    dev = rgb_coarse.device
    d_c = torch.empty_like(rgb_coarse)
    d_f = torch.empty_like(rgb_fine) if rgb_fine is not None else None
    d_b = torch.empty_like(beta) if beta is not None else None
    d_s = torch.empty_like(transient_sigmas) if (transient_sigmas is not None and beta is not None) else None
    n_ts = transient_sigmas.numel() if transient_sigmas is not None else 0
    gp = (_vp * 5)(*[(g.data_ptr() if g is not None else None) for g in g_terms])
    fn = _fn("nwb_nerfw_loss_bwd", [_vp] * 5 + [_i, _i64, C.c_float, C.c_float, C.POINTER(_vp * 5), _vp, _vp, _vp, _vp, _vp])
    with _timed("nwb_nerfw_loss_bwd", rgb_coarse.shape[0], dev=dev):
        L.check(fn(L.ptr(rgb_coarse), L.ptr(rgb_fine), L.ptr(beta), L.ptr(transient_sigmas), L.ptr(targets),
                   rgb_coarse.shape[0], n_ts, lambda_u, coef, C.byref(gp), L.ptr(d_c), L.ptr(d_f), L.ptr(d_b), L.ptr(d_s),
                   L.stream_ptr(dev)), "nwb_nerfw_loss_bwd")
    return d_c, d_f, d_b, d_s


# ---- metrics (f1), on-device ray generation (f3), eval output pack (f4) ---------------------------
def mse_psnr(pred: torch.Tensor, gt: torch.Tensor, clip01: bool = False) -> torch.Tensor:
    """-> (2,) = [mse, psnr] on the device (metrics.py:4-13), one launch."""
    pred, gt = _f32(pred), _f32(gt)
    if pred.shape != gt.shape:
        raise ValueError(f"mse_psnr: shapes differ {tuple(pred.shape)} vs {tuple(gt.shape)}")
    dev = pred.device
    out = torch.empty(2, device=dev, dtype=torch.float32)
    fn = _fn("nwb_mse_psnr", [_vp, _vp, _i64, _i, _vp, _vp, _vp])
    with _timed("nwb_mse_psnr", pred.numel(), dev=dev):
        L.check(fn(L.ptr(pred), L.ptr(gt), pred.numel(), int(clip01), L.ptr(_loss_scratch_for(dev)), L.ptr(out),
                   L.stream_ptr(dev)), "nwb_mse_psnr")
    return out


def gen_rays(H: int, W: int, K, c2w, near: float, far: float, first_pixel: int = 0, n_rays: int | None = None,
             device=None, out: torch.Tensor | None = None) -> torch.Tensor:
    """Rays (n,8) = [o, d, near, far] of pixels [first_pixel, first_pixel+n) of one view, row-major, made on the device
    (ray_utils.py:5-55 + blender.py:97-100).  K: (3,3) intrinsics, c2w: (3,4) pose -- small HOST arrays/tensors."""
    Kh = torch.as_tensor(K, dtype=torch.float64).cpu()
    pose = torch.as_tensor(c2w, dtype=torch.float32).cpu().contiguous()
    if pose.shape != (3, 4):
        raise ValueError(f"gen_rays: c2w must be (3,4), got {tuple(pose.shape)}")
    n = H * W - first_pixel if n_rays is None else int(n_rays)
    if out is None:
        out = torch.empty(n, 8, device=device or torch.device("cuda"), dtype=torch.float32)
    cbuf = (C.c_float * 12)(*pose.reshape(-1).tolist())
    fn = _fn("nwb_gen_rays", [_i, _i] + [C.c_float] * 4 + [C.POINTER(C.c_float * 12), C.c_float, C.c_float, _i64, _i64, _vp, _vp])
    with _timed("nwb_gen_rays", n, dev=out.device):
        L.check(fn(H, W, float(Kh[0, 0]), float(Kh[1, 1]), float(Kh[0, 2]), float(Kh[1, 2]), C.byref(cbuf), float(near),
                   float(far), int(first_pixel), n, L.ptr(out), L.stream_ptr(out.device)), "nwb_gen_rays")
    return out


def pack_rgb8(x: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """uint8(clip(x, 0, 1) * 255), same shape (eval.py:201-203), one launch."""
    x = _f32(x)
    if out is None:
        out = torch.empty(x.shape, device=x.device, dtype=torch.uint8)
    fn = _fn("nwb_pack_rgb8", [_vp, _i64, _vp, _vp])
    with _timed("nwb_pack_rgb8", x.numel(), dev=x.device):
        L.check(fn(L.ptr(x), x.numel(), L.ptr(out), L.stream_ptr(x.device)), "nwb_pack_rgb8")
    return out
