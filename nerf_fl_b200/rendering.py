"""Host-side mirror of the reference's ``models/rendering.py``: ``render_rays`` (:49-289).

Same signature, kwargs and result-dict keys; the body is a short schedule of
C-ABI calls into libnerfw_b200.so on torch's current stream:

    nwb_sample_coarse -> [nwb_ray_consts_fwd -> nwb_mlp_fwd -> nwb_composite_fwd]  (coarse)
    nwb_sample_pdf_merge -> [nwb_ray_consts_fwd -> nwb_mlp_fwd -> nwb_composite_fwd]  (fine)

and, under autograd, the mirrored backward (nwb_composite_bwd -> nwb_mlp_bwd -> nwb_ray_consts_bwd).
PyTorch allocates the buffers and orders the stream; nothing here computes on the host or falls
back to ATen math.  Per-sample inputs (encoded xyz, repeated dir/a/t rows, rendering.py:96-130) are
never materialised: points are generated and encoded inside the field kernel.
"""
from __future__ import annotations

import contextlib
import itertools

import os

import torch

from . import _lib as L
from . import kernels as K

__all__ = ["render_rays", "field_apply", "check_errors"]

_call_counter = itertools.count(1)
_err_flags: dict = {}


def _err_flag(dev) -> torch.Tensor:
    key = (dev.type, dev.index if dev.index is not None else torch.cuda.current_device())
    if key not in _err_flags:
        _err_flags[key] = torch.zeros(1, dtype=torch.int32, device=dev)
    return _err_flags[key]


def check_errors(device=None) -> None:
    """Synchronising check of the device-side error flag (embedding index out of range -- the
    reference's device-side assert of README.md:124).  render_rays itself never syncs."""
    for key, flag in list(_err_flags.items()):
        if device is not None and torch.device(device).index not in (None, key[1]):
            continue
        if int(flag.item()) != 0:
            flag.zero_()
            raise RuntimeError("nerf_fl_b200: embedding index out of range (ts outside [0, N_vocab))")


def _next_seed() -> int:
    return (torch.initial_seed() * 0x9E3779B97F4A7C15 + next(_call_counter)) & (2 ** 64 - 1)


def _param_list(model):
    sd = model.nwb_params()
    return list(sd.keys()), list(sd.values())


def _grad_sinks(names, params, dev):
    """Gradient sinks for one model: preassigned ``.grad`` views of a FlatGrads buffer are
    accumulated into in place (and reported to autograd as None); anything else gets zeros carved
    from one flat allocation that is handed to autograd."""
    direct = all(getattr(p, "_nwb_direct_grad", False) and p.grad is not None for p in params)
    if direct:
        return {n: p.grad for n, p in zip(names, params)}, True
    total = sum(p.numel() for p in params)
    flat = torch.zeros(total, device=dev, dtype=torch.float32)
    sinks, off = {}, 0
    for n, p in zip(names, params):
        sinks[n] = flat[off:off + p.numel()].view(p.shape)
        off += p.numel()
    return sinks, False


_BWD_STREAMS = os.environ.get("NWB_BWD_STREAMS", "1") != "0"
_FORK_MAX_TILES = int(os.environ.get("NWB_FORK_MAX_TILES", "8192"))     # fine-pass 128-point tiles
_side_streams = {}


def _side_stream(dev, which: int = 0) -> torch.cuda.Stream:
    key = (dev.index if dev.index is not None else torch.cuda.current_device(), which)
    if key not in _side_streams:
        _side_streams[key] = torch.cuda.Stream(device=dev)
    return _side_streams[key]


class _Pass:
    """Everything one inference() call (rendering.py:83-226) leaves behind for its backward."""
    __slots__ = ("model", "typ", "S", "transient", "mode", "desc", "packed", "feat", "raw", "z", "saved",
                 "noise", "keys", "use_a", "a_rows", "t_rows", "ws")


def _ray_consts(model, dirs, n_freq_dir, ts, emb_a, emb_t, a_rows, t_rows, transient):
    """Per-ray feature rows [dir | a | t] and the per-ray bias of dir_encoding / transient_encoding.0: everything of a
    pass that does not depend on the sample depths."""
    use_a = bool(model.encode_appearance)
    sd = {k: v.detach() for k, v in model.nwb_params().items()}
    feat, rb, _ = K.ray_consts_fwd(model.nwb_desc(), sd, dirs, n_freq_dir=n_freq_dir, ts=ts,
                                   emb_a=emb_a if (use_a and a_rows is None) else None,
                                   emb_t=emb_t if (transient and t_rows is None) else None,
                                   a_rows=a_rows if use_a else None, t_rows=t_rows if transient else None,
                                   transient=transient, err_flag=_err_flag(dirs.device))
    return feat, rb


def _run_pass(model, rays, z, *, typ, test_time, white_back, noise_std, noise, transient, dirs, n_freq_dir,
              n_freq_xyz, ts, emb_a, emb_t, a_rows, t_rows, save, consts=None):
    """One inference() call: field forward over the ray samples + volume rendering."""
    n, S = z.shape
    dev = rays.device
    P = _Pass()
    P.model, P.typ, P.S, P.transient, P.z, P.noise = model, typ, S, transient, z, noise
    P.desc = model.nwb_desc()
    P.packed = model.nwb_packed()
    sigma_only = typ == "coarse" and test_time                     # rendering.py:103-111
    P.use_a = bool(model.encode_appearance)
    P.a_rows, P.t_rows = a_rows, t_rows
    if P.desc.in_a > 0 and not P.use_a:
        raise RuntimeError("a 'coarse' NeRF built with encode_appearance=True cannot be evaluated "
                           "(its dir_encoding expects appearance columns the renderer never feeds, rendering.py:115-117)")
    P.mode = K.mlp_mode(sigma_only=sigma_only, transient=transient, save=save)
    rb = None
    P.feat = None
    if consts is not None:
        P.feat, rb = consts
    elif not sigma_only:
        P.feat, rb = _ray_consts(model, dirs, n_freq_dir, ts, emb_a, emb_t, a_rows, t_rows, transient)
    P.saved = K.aligned_empty(K.mlp_save_bytes(P.desc, n * S, P.mode), dev, 16) if save else None
    P.raw = K.mlp_fwd(P.desc, P.packed, n_rays=n, S=S, mode=P.mode, ray_bias=rb, rays=rays, z=z,
                      n_freq_xyz=n_freq_xyz, saved=P.saved)
    # the field backward's workspace exists from here on: its first word (max|d_raw|, filled by the compositing backward)
    # is cleared by the compositing forward, a whole pass ahead of its use
    P.ws = K.aligned_empty(K.mlp_bwd_workspace_bytes(P.desc, n * S, P.mode & ~L.NWB_MLP_SAVE), dev, 16) if save else None
    res = K.composite_fwd(P.raw.view(n, S, -1), z, typ=typ, transient=transient, white_back=white_back,
                          test_time=test_time, beta_min=float(model.beta_min), noise=noise, noise_std=noise_std,
                          clear_word=P.ws)
    P.keys = list(res.keys())
    return P, res


def _bwd_composite(P: _Pass, grads_by_key: dict, *, white_back, noise_std, seed=None):
    """First stage of _bwd_pass: cotangents of the maps -> d_raw (and max|d_raw| into the workspace header)."""
    n, S = P.z.shape
    return K.composite_bwd(P.raw.view(n, S, -1), P.z, grads_by_key, typ=P.typ, transient=P.transient,
                           white_back=white_back, noise=P.noise, noise_std=noise_std, seed=seed, amax_out=P.ws,
                           amax_cleared=True)


def _bwd_pass(P: _Pass, grads_by_key: dict, *, white_back, noise_std, ts, emb_a_w, emb_t_w, names, params, aux=None,
              seed=None, d_raw=None):
    """Backward of _run_pass -> (param sinks, direct?, d_emb_a, d_emb_t, d_a_rows, d_t_rows).  ``seed``: the NerfWLoss
    cotangents of this pass's maps are formed inside the compositing backward (K.composite_bwd)."""
    n, S = P.z.shape
    dev = P.z.device
    # the compositing backward also leaves max|d_raw| in the field backward's workspace header (the 8-bit images' scale):
    # no reduction launch between the two
    ws = P.ws
    if d_raw is None:
        d_raw = _bwd_composite(P, grads_by_key, white_back=white_back, noise_std=noise_std, seed=seed)
    bwd_mode = P.mode | L.NWB_MLP_DZ_AMAX_GIVEN
    sinks, direct = _grad_sinks(names, params, dev)
    d_rb = torch.zeros(n, 256, device=dev, dtype=torch.float32)
    # dgrad produces d_ray_bias and the dZ images; wgrad (dZ -> weight gradients) and the per-ray constants' backward
    # (d_ray_bias -> tail columns, biases, embedding rows) are independent of each other: with `aux` (a stream, graph
    # capture only) the second runs next to the first
    K.mlp_bwd(P.desc, P.packed, P.raw, d_raw.view(n * S, -1), P.saved, n_rays=n, S=S, mode=bwd_mode, grads=sinks,
              d_ray_bias=d_rb, workspace=ws, phases=("dgrad",))
    if aux is not None:
        aux.wait_stream(torch.cuda.current_stream(dev))
    K.mlp_bwd(P.desc, P.packed, P.raw, d_raw.view(n * S, -1), P.saved, n_rays=n, S=S, mode=bwd_mode, grads=sinks,
              d_ray_bias=d_rb, workspace=ws, phases=("wgrad",))
    sd = {k: v.detach() for k, v in zip(names, params)}
    d_a_rows = torch.empty_like(P.a_rows) if (P.use_a and P.a_rows is not None) else None
    d_t_rows = torch.empty_like(P.t_rows) if (P.transient and P.t_rows is not None) else None
    d_ea = d_et = None
    ea_direct = et_direct = False
    if P.use_a and P.a_rows is None and emb_a_w is not None and emb_a_w.requires_grad:
        ea_direct = getattr(emb_a_w, "_nwb_direct_grad", False) and emb_a_w.grad is not None
        d_ea = emb_a_w.grad if ea_direct else torch.zeros_like(emb_a_w)
    if P.transient and P.t_rows is None and emb_t_w is not None and emb_t_w.requires_grad:
        et_direct = getattr(emb_t_w, "_nwb_direct_grad", False) and emb_t_w.grad is not None
        d_et = emb_t_w.grad if et_direct else torch.zeros_like(emb_t_w)
    n_vocab = emb_a_w.shape[0] if emb_a_w is not None else (emb_t_w.shape[0] if emb_t_w is not None else 0)
    with (torch.cuda.stream(aux) if aux is not None else contextlib.nullcontext()):
        K.ray_consts_bwd(P.desc, sd, P.feat, d_rb, ts=ts, n_vocab=n_vocab, transient=P.transient, grads=sinks,
                         d_emb_a=d_ea, d_emb_t=d_et, d_a_rows=d_a_rows, d_t_rows=d_t_rows)
    if aux is not None:
        torch.cuda.current_stream(dev).wait_stream(aux)
    return sinks, direct, (None if ea_direct else d_ea), (None if et_direct else d_et), d_a_rows, d_t_rows


_LOSS_KEYS = ("loss_c_l", "loss_f_l", "loss_b_l", "loss_s_l", "loss_total")


def _detached(t):
    return t.detach() if t is not None else None


class _RenderFn(torch.autograd.Function):
    """The whole of render_rays as one autograd node (coarse pass, resampling, fine pass)."""

    @staticmethod
    def forward(ctx, cfg, rays, ts, view_dir, a_embedded, t_embedded, emb_a_w, emb_t_w, n_c, *params):
        models = cfg["models"]
        coarse, fine = models["coarse"], models.get("fine")
        rng = cfg["rng"]
        n = rays.shape[0]
        Ns, Ni = cfg["N_samples"], cfg["N_importance"]
        perturb, noise_std, test_time = cfg["perturb"], cfg["noise_std"], cfg["test_time"]
        save = cfg["save"]
        seed = cfg["seed"]
        dirs = view_dir if view_dir is not None else rays[:, 3:6].contiguous()                # rendering.py:238

        def sigma_noise(shape, key):                                                          # :151
            if noise_std == 0.0:
                return None
            t = rng.get(key)
            return t if t is not None else torch.randn(shape, device=rays.device, dtype=torch.float32)

        common = dict(test_time=test_time, white_back=cfg["white_back"], noise_std=noise_std, dirs=dirs,
                      n_freq_dir=cfg["n_freq_dir"], n_freq_xyz=cfg["n_freq_xyz"], save=save)
        # the fine pass's per-ray constants do not depend on the coarse pass: inside a graph capture they go on a side
        # branch next to it
        transient_f = bool(Ni > 0 and cfg["output_transient"] and fine.encode_transient)       # :281
        consts_c = consts_f = None
        fork = Ni > 0 and _BWD_STREAMS and rays.is_cuda and torch.cuda.is_current_stream_capturing()
        if fork:
            cur, side = torch.cuda.current_stream(rays.device), _side_stream(rays.device)
            side.wait_stream(cur)
            with torch.cuda.stream(side):
                consts_f = _ray_consts(fine, dirs, cfg["n_freq_dir"], ts, emb_a_w, emb_t_w, a_embedded, t_embedded,
                                       transient_f)
            if not test_time:
                # the coarse pass's constants do not depend on the sample depths either: next to the depth sampling
                side_c = _side_stream(rays.device, 5)
                side_c.wait_stream(cur)
                with torch.cuda.stream(side_c):
                    consts_c = _ray_consts(coarse, dirs, cfg["n_freq_dir"], None, None, None, None, None, False)
        z = K.sample_coarse(rays, Ns, cfg["use_disp"], perturb, rng.get("perturb_rand"), seed)   # :244-259
        if consts_c is not None:
            cur.wait_stream(side_c)
        Pc, res = _run_pass(coarse, rays, z, typ="coarse", noise=sigma_noise((n, Ns), "noise_coarse"),
                            transient=False, ts=None, emb_a=None, emb_t=None, a_rows=None, t_rows=None, consts=consts_c,
                            **common)
        if fork:
            cur.wait_stream(side)
        Pf = None
        if Ni > 0:
            z_f = K.sample_pdf_merge(z, res["weights_coarse"], Ni, det=(perturb == 0),              # :267-272
                                     u=rng.get("u") if perturb > 0 else None, seed=seed)
            transient = transient_f
            noise_f = None if transient else sigma_noise((n, Ns + Ni), "noise_fine")
            Pf, res_f = _run_pass(fine, rays, z_f, typ="fine", noise=noise_f, transient=transient, ts=ts,
                                  emb_a=emb_a_w, emb_t=emb_t_w, a_rows=a_embedded, t_rows=t_embedded, consts=consts_f,
                                  **common)
            res.update(res_f)
        # NerfWLoss over the maps, inside this node (cfg["loss"], set by render_rays(loss=...)): the five terms become
        # extra outputs and the backward forms their cotangents in the compositing kernels (no loss-backward launch, no
        # cotangent maps).  The forward launch is off the critical path when the caller promises a backward (TrainStep).
        ctx.loss = None
        spec = cfg.get("loss")
        if spec is not None and save:
            beta = res.get("beta")
            tsig = res.get("transient_sigmas") if beta is not None else None
            lstream, deferred = None, None
            loss_args = (res["rgb_coarse"].detach(), _detached(res.get("rgb_fine")), _detached(beta), _detached(tsig),
                         spec["targets"], spec["lambda_u"], spec["coef"])
            if fork and spec.get("join_in_backward"):
                # the caller promises a backward inside the same capture (TrainStep): the terms are only read after the
                # step, so their kernel is launched from the backward, on a branch of its own behind the fine pass's
                # compositing backward (next to it, at the head of the backward's critical path, it was in the way)
                lstream = _side_stream(rays.device, 3)
                terms = torch.empty(5, device=rays.device, dtype=torch.float32)
                deferred = lambda: K.nerfw_loss_fwd(*loss_args, out=terms)
            else:
                terms = K.nerfw_loss_fwd(*loss_args)
            ctx.loss = dict(rgb_c=res["rgb_coarse"].detach(), rgb_f=_detached(res.get("rgb_fine")), beta=_detached(beta),
                            n_ts=tsig.numel() if tsig is not None else 0, targets=spec["targets"],
                            coef=spec["coef"], lambda_u=spec["lambda_u"], stream=lstream, deferred=deferred)
            for i, k in enumerate(_LOSS_KEYS):
                res[k] = terms[i]
        cfg["keys"] = list(res.keys())
        ctx.cfg, ctx.Pc, ctx.Pf, ctx.n_c = cfg, Pc, Pf, n_c
        ctx.ts, ctx.emb_a_w, ctx.emb_t_w = ts, emb_a_w, emb_t_w
        ctx.params = params
        # the loss touches 4 of the ~14 outputs (losses.py:37-45); the others' cotangents stay None instead of being
        # materialised as zero tensors (one fill launch each) and the compositing backward skips their terms
        ctx.set_materialize_grads(False)
        outs = tuple(res.values())
        if not save:
            ctx.mark_non_differentiable(*outs)
        return outs

    @staticmethod
    def backward(ctx, *gouts):
        cfg, Pc, Pf, n_c = ctx.cfg, ctx.Pc, ctx.Pf, ctx.n_c
        g = {k: (v.contiguous() if v is not None else None) for k, v in zip(cfg["keys"], gouts)}
        names_c, names_f = cfg["names_c"], cfg["names_f"]
        params_c, params_f = ctx.params[:n_c], ctx.params[n_c:]
        kw = dict(white_back=cfg["white_back"], noise_std=cfg["noise_std"])
        seed_c = seed_f = None
        if ctx.loss is not None:
            lo = ctx.loss
            g_terms = [g.pop(k, None) for k in _LOSS_KEYS]
            if any(t is not None for t in g_terms):
                common = dict(targets=lo["targets"], g_terms=g_terms, coef=lo["coef"], lambda_u=lo["lambda_u"],
                              n_rays=Pc.z.shape[0])
                seed_c = dict(rgb=lo["rgb_c"], beta=None, n_ts=0, fine=False, **common)
                if lo["rgb_f"] is not None:
                    seed_f = dict(rgb=lo["rgb_f"], beta=lo["beta"], n_ts=lo["n_ts"], fine=True, **common)
        out_params = []
        d_ea = d_et = d_a = d_t = None
        # The coarse and the fine chain (compositing bwd -> dgrad -> wgrad -> per-ray constants) share nothing but the
        # cotangent dict: with both models' gradients accumulated in place (FlatTrainState) the coarse chain runs on a
        # side stream, so each persistent kernel's last partial wave of tiles is filled by the other chain's CTAs.
        # Only inside a CUDA-graph capture (TrainStep): replayed, the two branches cost nothing on the host, while eager
        # launches are host-bound and the extra stream switches only add to that ([B200] 1.93 vs 1.64 ms eager).
        fork = (Pf is not None and _BWD_STREAMS and Pc.z.is_cuda and torch.cuda.is_current_stream_capturing()
                and all(getattr(p, "_nwb_direct_grad", False) and p.grad is not None for p in ctx.params))
        # Two whole chains side by side only pay while a kernel's last partial wave is a visible share of it ([B200]
        # 1024 rays 64+64: +6 %, 4096: +1.5 %; 16384 rays 128+128: -7 %, long kernels of different kinds sharing SMs
        # and L2): above _FORK_MAX_TILES fine tiles only the small side branches remain.
        fork_chains = fork and Pf.z.numel() <= _FORK_MAX_TILES * 128
        d_raw_f = None
        if fork_chains:
            cur = torch.cuda.current_stream(Pc.z.device)
            side = _side_stream(Pc.z.device)
            # the fine pass's compositing backward goes first and alone: it heads the longer chain, and next to the coarse
            # pass's it takes twice as long ([B200, tools/graph_timeline.py] 18 us against 8)
            d_raw_f = _bwd_composite(Pf, g, seed=seed_f, **kw)
            side.wait_stream(cur)
            if ctx.loss is not None and ctx.loss.get("deferred") is not None:
                ctx.loss["stream"].wait_stream(cur)
                with torch.cuda.stream(ctx.loss["stream"]):
                    ctx.loss["deferred"]()
                ctx.loss["deferred"] = None
            with torch.cuda.stream(side):
                sinks_c, direct_c, *_ = _bwd_pass(Pc, g, ts=None, emb_a_w=None, emb_t_w=None, names=names_c,
                                                  params=params_c, aux=_side_stream(Pc.z.device, 1), seed=seed_c, **kw)
                # the coarse model's gradients are complete here while the fine chain is still running: a caller that owns
                # the optimizer (TrainStep) updates the coarse model now, on this branch
                if direct_c and cfg.get("after_coarse_chain") is not None:
                    cfg["after_coarse_chain"]()
        else:
            sinks_c, direct_c, *_ = _bwd_pass(Pc, g, ts=None, emb_a_w=None, emb_t_w=None, names=names_c,
                                              params=params_c, aux=_side_stream(Pc.z.device, 1) if fork else None,
                                              seed=seed_c, **kw)
        out_params += [None if direct_c else sinks_c[k] for k in names_c]
        if Pf is not None:
            sinks_f, direct_f, d_ea, d_et, d_a, d_t = _bwd_pass(Pf, g, ts=ctx.ts, emb_a_w=ctx.emb_a_w,
                                                                emb_t_w=ctx.emb_t_w, names=names_f, params=params_f,
                                                                aux=_side_stream(Pc.z.device, 2) if fork else None,
                                                                seed=seed_f, d_raw=d_raw_f, **kw)
            out_params += [None if direct_f else sinks_f[k] for k in names_f]
            # the fine chain was started first and holds the SMs first: its gradients are complete here, while the coarse
            # chain's last kernels still run on the other branch -- a caller that owns the optimizer updates the fine model now
            if fork_chains and direct_f and cfg.get("after_fine_chain") is not None:
                cfg["after_fine_chain"]()
        if fork_chains:
            cur.wait_stream(side)
        if ctx.loss is not None and ctx.loss["stream"] is not None:
            if ctx.loss.get("deferred") is not None:         # no forked chains in this backward: launch it here
                ctx.loss["deferred"]()
                ctx.loss["deferred"] = None
            else:
                torch.cuda.current_stream(Pc.z.device).wait_stream(ctx.loss["stream"])      # the loss value joins here
        # inputs: cfg, rays, ts, view_dir, a_embedded, t_embedded, emb_a_w, emb_t_w, n_c, *params
        return (None, None, None, None, d_a, d_t, d_ea, d_et, None, *out_params)


def render_rays(models, embeddings, rays, ts, N_samples=64, use_disp=False, perturb=0, noise_std=1,
                N_importance=0, chunk=1024 * 32, white_back=False, test_time=False, **kwargs):
    """Drop-in for the reference's render_rays (models/rendering.py:49-62; semantics :228-289).

    ``chunk`` is accepted for compatibility and ignored: it only bounds activation memory in the
    reference (rendering.py:101-130), and activations stay on chip here.  Extra kwargs understood
    (all optional): the reference's ``a_embedded``, ``t_embedded``, ``output_transient``,
    ``view_dir``, ``current_epoch`` (ignored: BARF only); plus ``rng`` -- a dict of pre-drawn
    tensors ``perturb_rand`` (N,Ns) U[0,1), ``noise_coarse`` (N,Ns) N(0,1), ``u`` (N,Ni) U[0,1),
    ``noise_fine`` (N,Ns+Ni) N(0,1) replacing the draws of rendering.py:257,151,30 (parity runs) --
    and ``seed`` for the in-kernel Philox streams used otherwise; ``loss`` -- ``dict(fn=NerfWLoss(...), targets=rgbs)``:
    the objective (losses.py:35-50) is evaluated inside the same autograd node and comes back as
    ``result['loss_terms']`` (the dict NerfWLoss returns, ``.total`` = their sum); its backward is folded into the
    compositing kernels, so a training step launches neither a loss backward nor cotangent maps (TrainStep uses it).
    Gradients flow to every NeRF parameter, to ``embeddings['a'|'t'].weight`` and to explicit
    ``a_embedded`` / ``t_embedded`` rows; not to rays / view_dir (no pose refinement on this path).
    With ``test_time=True`` outputs carry no gradient (eval.py:80 runs under no_grad)."""
    if not rays.is_cuda:
        raise RuntimeError("nerf_fl_b200.render_rays needs CUDA tensors: there is no CPU path")
    coarse = models["coarse"]
    fine = models.get("fine") if N_importance > 0 else None
    if N_importance > 0 and fine is None:
        raise KeyError("fine")                                        # the reference fails the same way (:274)
    rays_c = K._f32(rays)
    if rays_c.shape[1] < 8:
        raise RuntimeError("rays must be (N, >=8): [o, d, near, far] (rendering.py:232-233)")
    n_freq_xyz = (coarse.in_channels_xyz - 3) // 6
    n_freq_dir = (coarse.in_channels_dir - 3) // 6
    for key, width in (("xyz", coarse.in_channels_xyz), ("dir", coarse.in_channels_dir)):
        e = embeddings.get(key)
        nf = getattr(e, "N_freqs", None)
        if nf is not None and 6 * nf + 3 != width:
            raise RuntimeError(f"embeddings['{key}'] yields {6 * nf + 3} channels but the model expects {width}")
    emb_a_w = emb_t_w = None
    a_embedded, t_embedded = kwargs.get("a_embedded"), kwargs.get("t_embedded")
    if fine is not None:
        if fine.encode_appearance and a_embedded is None:
            emb_a_w = embeddings["a"].weight
        if fine.encode_transient and kwargs.get("output_transient", True) and t_embedded is None:
            emb_t_w = embeddings["t"].weight
        if (emb_a_w is not None or emb_t_w is not None) and ts is None:
            raise RuntimeError("ts is required when appearance / transient embeddings are looked up")
    if ts is not None:
        ts = ts.to(torch.int64).contiguous()
    n = rays_c.shape[0]
    if a_embedded is not None:
        a_embedded = K._f32(a_embedded).expand(n, -1).contiguous()
    if t_embedded is not None:
        t_embedded = K._f32(t_embedded).expand(n, -1).contiguous()
    view_dir = kwargs.get("view_dir")
    if view_dir is not None:
        view_dir = K._f32(view_dir)
    names_c, params_c = _param_list(coarse)
    names_f, params_f = _param_list(fine) if fine is not None else ([], [])
    diff_inputs = params_c + params_f + [t for t in (emb_a_w, emb_t_w, a_embedded, t_embedded) if t is not None]
    save = (not test_time) and torch.is_grad_enabled() and any(t.requires_grad for t in diff_inputs)
    cfg = dict(models={"coarse": coarse, "fine": fine}, N_samples=int(N_samples), N_importance=int(N_importance),
               use_disp=bool(use_disp), perturb=float(perturb), noise_std=float(noise_std),
               white_back=bool(white_back), test_time=bool(test_time),
               output_transient=bool(kwargs.get("output_transient", True)), n_freq_xyz=n_freq_xyz,
               n_freq_dir=n_freq_dir, rng=dict(kwargs.get("rng") or {}), save=save,
               seed=int(kwargs["seed"]) if "seed" in kwargs else _next_seed(), names_c=names_c, names_f=names_f)
    for k, v in cfg["rng"].items():
        cfg["rng"][k] = K._f32(v)
    if n == 0:
        raise RuntimeError("render_rays called with zero rays")
    for hook in ("after_coarse_chain", "after_fine_chain"):
        if kwargs.get(hook) is not None:
            cfg[hook] = kwargs[hook]
    loss = kwargs.get("loss")
    if loss is not None:
        fn, targets = loss["fn"], K._f32(loss["targets"])
        if targets.shape != (n, 3):
            raise RuntimeError(f"loss targets must be ({n}, 3), got {tuple(targets.shape)}")
        cfg["loss"] = dict(targets=targets, coef=float(fn.coef), lambda_u=float(fn.lambda_u),
                           join_in_backward=bool(loss.get("join_in_backward")))
    outs = _RenderFn.apply(cfg, rays_c, ts, view_dir, a_embedded, t_embedded, emb_a_w, emb_t_w,
                           len(params_c), *params_c, *params_f)
    res = dict(zip(cfg["keys"], outs))
    if _LOSS_KEYS[0] in res:
        from .objective import LossTerms
        c_l, f_l, b_l, s_l, total = (res.pop(k) for k in _LOSS_KEYS)
        terms = LossTerms(c_l=c_l)
        if "rgb_fine" in res:
            terms["f_l"] = f_l
        if "beta" in res:
            terms["b_l"], terms["s_l"] = b_l, s_l
        terms.total = total
        res["loss_terms"] = terms
    return res


# ---- NeRF.forward on explicit embedded rows (models/nerf.py:153-212) ---------------------------------
class _FieldFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, model, x, sigma_only, transient, save, names, *params):
        B = x.shape[0]
        dev = x.device
        desc, packed = model.nwb_desc(), model.nwb_packed()
        ix, idr = model.in_channels_xyz, model.in_channels_dir
        ia = model.in_channels_a if model.encode_appearance else 0
        sd = {k: v.detach() for k, v in zip(names, params)}
        mode = K.mlp_mode(sigma_only=sigma_only, transient=transient, save=save)
        feat = rb = a_rows = t_rows = None
        if not sigma_only:
            if desc.in_a > 0 and ia == 0:
                raise RuntimeError("a 'coarse' NeRF built with encode_appearance=True cannot be evaluated")
            a_rows = x[:, ix + idr:ix + idr + ia].contiguous() if ia else None
            t_rows = x[:, ix + idr + ia:ix + idr + ia + model.in_channels_t].contiguous() if transient else None
            feat, rb, _ = K.ray_consts_fwd(desc, sd, x[:, ix:ix + idr].contiguous(), dir_is_encoded=True,
                                           a_rows=a_rows, t_rows=t_rows, transient=transient, err_flag=_err_flag(dev))
        saved = K.aligned_empty(K.mlp_save_bytes(desc, B, mode), dev, 16) if save else None
        raw = K.mlp_fwd(desc, packed, n_rays=B, S=1, mode=mode, ray_bias=rb, x_enc=x[:, :ix].contiguous(), saved=saved)
        ctx.model, ctx.names, ctx.params, ctx.mode, ctx.transient = model, names, params, mode, transient
        ctx.desc, ctx.packed, ctx.feat, ctx.saved, ctx.raw, ctx.xshape = desc, packed, feat, saved, raw, x.shape
        ctx.ia = ia
        if not save:
            ctx.mark_non_differentiable(raw)
        return raw

    @staticmethod
    def backward(ctx, d_raw):
        model, names, params = ctx.model, ctx.names, ctx.params
        B = ctx.xshape[0]
        dev = d_raw.device
        sinks, direct = _grad_sinks(names, params, dev)
        d_rb = torch.zeros(B, 256, device=dev, dtype=torch.float32)
        K.mlp_bwd(ctx.desc, ctx.packed, ctx.raw, d_raw.contiguous(), ctx.saved, n_rays=B, S=1, mode=ctx.mode,
                  grads=sinks, d_ray_bias=d_rb)
        ix, idr, ia = model.in_channels_xyz, model.in_channels_dir, ctx.ia
        d_a = torch.empty(B, ia, device=dev) if ia else None
        d_t = torch.empty(B, model.in_channels_t, device=dev) if ctx.transient else None
        K.ray_consts_bwd(ctx.desc, {k: v.detach() for k, v in zip(names, params)}, ctx.feat, d_rb, transient=ctx.transient,
                         grads=sinks, d_a_rows=d_a, d_t_rows=d_t)
        dx = torch.zeros(ctx.xshape, device=dev, dtype=torch.float32)   # no gradient to the xyz / dir columns
        if d_a is not None:
            dx[:, ix + idr:ix + idr + ia] = d_a
        if d_t is not None:
            dx[:, ix + idr + ia:ix + idr + ia + model.in_channels_t] = d_t
        return (None, dx, None, None, None, None, *[None if direct else sinks[k] for k in names])


def field_apply(model, x, sigma_only=False, output_transient=True):
    """NeRF.forward: x (B, in_xyz [+ in_dir + in_a [+ in_t]]) -> (B, 1 | 4 | 9)."""
    if not x.is_cuda:
        raise RuntimeError("nerf_fl_b200.NeRF.forward needs CUDA tensors: there is no CPU path")
    x = K._f32(x)
    ix, idr = model.in_channels_xyz, model.in_channels_dir
    ia = model.in_channels_a if model.encode_appearance else 0
    transient = bool(output_transient) and not sigma_only
    if transient and not model.encode_transient:
        raise AttributeError("NeRF has no transient head (nerf.py:200 would fail on transient_encoding)")
    want = ix if sigma_only else ix + idr + ia + (model.in_channels_t if transient else 0)
    if x.dim() != 2 or x.shape[1] != want:
        raise RuntimeError(f"NeRF.forward expects (B, {want}) rows, got {tuple(x.shape)} (torch.split at nerf.py:171-181)")
    if x.shape[0] == 0:
        return x.new_zeros(0, 1 if sigma_only else (9 if transient else 4))
    names, params = _param_list(model)
    save = (not sigma_only) and torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in params))
    return _FieldFn.apply(model, x, sigma_only, transient, save, names, *params)
