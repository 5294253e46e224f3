"""in_channels_xyz = 93 (N_emb_xyz = 15, the reference's phototourism notebook shape: test_phototourism.ipynb cell 1,
models/nerf.py:83,121-128, eval.py:126,139-143) through the oracle (CPU) and through the product (GPU): field forward in
the four head modes, the notebook's test_time render at 128+128 with N_vocab 1500, and a training step with gradients.
Golden vectors: tests/golden/xyz15.npz, written by oracle/make_golden_xyz15.py from the unmodified reference."""
import pytest
import torch

from nerf_fl_b200 import synthetic as syn
from oracle import nerfw_oracle as O

IX, NV = 93, 1500
DEV = "cuda"


def maxerr(a, b):
    return (a.detach().double().cpu() - b.detach().double().cpu()).abs().max().item()


def sub(g, tag):
    return {k[len(tag) + 1:]: v for k, v in g.items() if k.startswith(tag + "/") and "/grad" not in k and not k.endswith("/loss")}


def oracle_params():
    return (syn.make_field_params(41, "coarse", in_xyz=IX), syn.make_field_params(42, "fine", True, True, in_xyz=IX),
            syn.make_embedding(31, NV, 48), syn.make_embedding(32, NV, 16))


# ---- CPU: the oracle against the reference's vectors ------------------------------------------------------------------
def test_oracle_field_forward_xyz15(golden):
    g = golden("xyz15")
    x = g["field/x"]
    pc = syn.make_field_params(21, "coarse", in_xyz=IX)
    pa = syn.make_field_params(22, "fine", True, False, in_xyz=IX)
    pt = syn.make_field_params(23, "fine", True, True, in_xyz=IX)
    assert maxerr(O.field_forward(pc, x[:, :IX], in_xyz=IX, sigma_only=True), g["field/sigma_only"]) <= 1e-6
    assert maxerr(O.field_forward(pc, x[:, :IX + 27], in_xyz=IX, output_transient=False), g["field/plain"]) <= 1e-6
    assert maxerr(O.field_forward(pa, x[:, :IX + 75], in_xyz=IX, in_dir_a=75, output_transient=False), g["field/a"]) <= 1e-6
    assert maxerr(O.field_forward(pt, x, in_xyz=IX, in_dir_a=75, output_transient=True), g["field/at"]) <= 1e-6


def test_oracle_render_rays_xyz15(golden):
    g = golden("xyz15")
    pc, pf, ea, et = oracle_params()
    rng = {k[4:]: v for k, v in g.items() if k.startswith("rng_")}
    out = O.render_rays(pc, pf, ea, et, g["rays"], g["ts"], n_samples=128, n_importance=128, perturb=0, noise_std=0,
                        test_time=True, n_freq_xyz=15, encode_a=True, encode_t=True, beta_min=0.03, rng=rng)
    want = sub(g, "test")
    assert set(out) == set(want)
    for k, w in want.items():
        assert maxerr(out[k], w) <= 2e-5 * max(1.0, w.abs().max().item()), k
    for v in (pc, pf):
        for t in v.values():
            t.requires_grad_(True)
    out = O.render_rays(pc, pf, ea, et, g["rays"], g["ts"], n_samples=128, n_importance=128, perturb=1.0, noise_std=0.0,
                        n_freq_xyz=15, encode_a=True, encode_t=True, beta_min=0.03, rng=rng)
    loss = sum(O.nerfw_loss(out, g["targets"]).values())
    assert abs(loss.item() - g["train/loss"].item()) <= 1e-5
    names = [("coarse." + k, v) for k, v in pc.items()] + [("fine." + k, v) for k, v in pf.items()]
    for (k, _), gr in zip(names, torch.autograd.grad(loss, [v for _, v in names])):
        assert abs(gr.norm().item() - g[f"train/grad_norm/{k}"].item()) <= 1e-4 * max(g[f"train/grad_norm/{k}"].item(), 1e-6), k


# ---- GPU: the product against the same vectors ---------------------------------------------------------------------
def _models():
    from nerf_fl_b200.nerf import NeRF, PosEmbedding
    c = NeRF("coarse", in_channels_xyz=IX)
    c.load_state_dict(syn.make_field_params(41, "coarse", in_xyz=IX))
    f = NeRF("fine", in_channels_xyz=IX, encode_appearance=True, encode_transient=True, beta_min=0.03)
    f.load_state_dict(syn.make_field_params(42, "fine", True, True, in_xyz=IX))
    emb = {"xyz": PosEmbedding(14, 15), "dir": PosEmbedding(3, 4),
           "a": torch.nn.Embedding.from_pretrained(syn.make_embedding(31, NV, 48), freeze=False).to(DEV),
           "t": torch.nn.Embedding.from_pretrained(syn.make_embedding(32, NV, 16), freeze=False).to(DEV)}
    return {"coarse": c.to(DEV), "fine": f.to(DEV)}, emb


@pytest.mark.gpu
@pytest.mark.parametrize("head", ["sigma_only", "plain", "a", "at"])
def test_field_forward_xyz15_golden(golden, head):
    """explicit encoded rows (the NeRF.forward surface), both 64-column xyz blocks in use"""
    from nerf_fl_b200 import kernels as K
    g = golden("xyz15")
    x = g["field/x"].to(DEV)
    enc_a, enc_t = head in ("a", "at"), head == "at"
    seed = {"sigma_only": 21, "plain": 21, "a": 22, "at": 23}[head]
    sd = {k: v.to(DEV) for k, v in syn.make_field_params(seed, "coarse" if seed == 21 else "fine", enc_a, enc_t, in_xyz=IX).items()}
    desc = K.field_desc(IX, 27, 48 if enc_a else 0, 16 if enc_t else 0)
    packed = K.field_pack(desc, sd)
    rb = None
    if head != "sigma_only":
        _, rb, _ = K.ray_consts_fwd(desc, sd, x[:, IX:IX + 27].contiguous(), dir_is_encoded=True,
                                    a_rows=x[:, IX + 27:IX + 75].contiguous() if enc_a else None,
                                    t_rows=x[:, IX + 75:].contiguous() if enc_t else None, transient=enc_t)
    got = K.mlp_fwd(desc, packed, n_rays=x.shape[0], S=1, mode=K.mlp_mode(sigma_only=head == "sigma_only", transient=enc_t),
                    ray_bias=rb, x_enc=x[:, :IX].contiguous())
    want = g[f"field/{head}"]
    assert got.shape == want.shape
    assert maxerr(got, want) <= 3e-3 * max(1.0, want.abs().max().item())


@pytest.mark.gpu
def test_render_rays_xyz15_notebook_test_time(golden):
    from nerf_fl_b200.rendering import render_rays, check_errors
    g = golden("xyz15")
    models, emb = _models()
    rng = {k[4:]: v.to(DEV) for k, v in g.items() if k.startswith("rng_")}
    with torch.no_grad():
        out = render_rays(models, emb, g["rays"].to(DEV), g["ts"].to(DEV), 128, False, 0, 0, 128, 32768, False, test_time=True, rng=rng)
    check_errors()
    want = sub(g, "test")
    assert set(out) == set(want)
    for k, w in want.items():
        lim = 1e-3 * max(1.0, w.abs().max().item()) if k == "transient_sigmas" else 1e-3
        assert maxerr(out[k], w) <= lim, (k, maxerr(out[k], w))


@pytest.mark.gpu
def test_render_rays_xyz15_train_step_gradients(golden):
    """forward keys within 1e-3 absolute of the reference; gradient of each network within 1e-2 relative of the fp32
    oracle (whose own gradients test_oracle_render_rays_xyz15 pins to the reference's), first-layer tensors (the 93-column
    operands, both xyz blocks) against the reference's stored tensors"""
    from nerf_fl_b200.rendering import render_rays, check_errors
    g = golden("xyz15")
    models, emb = _models()
    rng = {k[4:]: v.to(DEV) for k, v in g.items() if k.startswith("rng_")}
    out = render_rays(models, emb, g["rays"].to(DEV), g["ts"].to(DEV), 128, False, 1.0, 0.0, 128, 32768, False, rng=rng)
    check_errors()
    want = sub(g, "train")
    assert set(out) == set(want)
    for k, w in want.items():
        lim = 1e-3 * max(1.0, w.abs().max().item()) if k == "transient_sigmas" else 1e-3
        assert maxerr(out[k], w) <= lim, (k, maxerr(out[k], w))
    loss = sum(O.nerfw_loss(out, g["targets"].to(DEV)).values())
    assert abs(loss.item() - g["train/loss"].item()) <= 1e-4 * abs(g["train/loss"].item())
    named = [(f"coarse.{k}", p) for k, p in models["coarse"].named_parameters()] + [(f"fine.{k}", p) for k, p in models["fine"].named_parameters()]
    got = {k: gr.detach().double().cpu() for (k, _), gr in zip(named, torch.autograd.grad(loss, [p for _, p in named]))}
    def oracle_grads(model):
        pc, pf, ea, et = oracle_params()
        onamed = [("coarse." + k, v) for k, v in pc.items()] + [("fine." + k, v) for k, v in pf.items()]
        for _, v in onamed:
            v.requires_grad_(True)
        oout = O.render_rays(pc, pf, ea, et, g["rays"], g["ts"], n_samples=128, n_importance=128, perturb=1.0, noise_std=0.0,
                             n_freq_xyz=15, encode_a=True, encode_t=True, beta_min=0.03, rng={k: v.cpu() for k, v in rng.items()},
                             bf16_operands=model, wgrad8=model)
        oloss = sum(O.nerfw_loss(oout, g["targets"]).values())
        return {k: gr.double() for (k, _), gr in zip(onamed, torch.autograd.grad(oloss, [v for _, v in onamed]))}
    want_g, model_g = oracle_grads(False), oracle_grads(True)

    def rel(a, b, keys):
        num = sum(((a[k] - b[k]).norm() ** 2).item() for k in keys)
        return (num / max(sum((b[k].norm() ** 2).item() for k in keys), 1e-300)) ** 0.5
    for net in ("coarse.", "fine."):
        keys = [k for k in want_g if k.startswith(net)]
        assert rel(got, want_g, keys) <= 1e-2, (net, rel(got, want_g, keys))
    # the layers fed by the 93-column operand (both xyz blocks; the skip layer's weight is 256 x (93 + 256)): per tensor
    # against the oracle's model of the kernels' operand rounding, and no further from fp32 than that model itself is
    for k in ("fine.xyz_encoding_5.0.weight", "fine.xyz_encoding_1.0.weight", "coarse.xyz_encoding_5.0.weight", "coarse.xyz_encoding_1.0.weight"):
        assert got[k].shape == want_g[k].shape
        assert rel(got, model_g, [k]) <= 3e-2, (k, rel(got, model_g, [k]))
        assert rel(got, want_g, [k]) <= 1.5 * rel(model_g, want_g, [k]) + 1e-2, (k, rel(got, want_g, [k]), rel(model_g, want_g, [k]))
    for k in ("fine.xyz_encoding_1.0.weight", "coarse.xyz_encoding_1.0.weight"):
        ref = g[f"train/grad/{k}"].double()
        assert (want_g[k] - ref).norm() <= 1e-4 * ref.norm()            # oracle == reference on the stored tensor
