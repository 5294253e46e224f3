"""Per-tensor gradient report of one NeRF-W training step: product (CUDA) vs the oracle's fp32 path and
vs its bf16-operand model (oracle/nerfw_oracle.py), on the golden inputs or a seeded larger batch."""
import math, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_fl_b200 import synthetic as syn
from nerf_fl_b200.nerf import NeRF, PosEmbedding
from nerf_fl_b200.rendering import render_rays
from oracle import nerfw_oracle as O

DEV = "cuda"


def inputs(n):
    rays = syn.blender_rays(n, seed=5)
    gen = torch.Generator().manual_seed(7)
    ts = torch.randint(0, 100, (n,), generator=gen)
    tgt = torch.rand(n, 3, generator=gen)
    rng = {"perturb_rand": torch.rand(n, 64, generator=gen), "u": torch.rand(n, 64, generator=gen)}
    return rays, ts, tgt, rng


def oracle(rays, ts, tgt, rng, scale, bf):
    ea, et = syn.make_embedding(31, 100, 48), syn.make_embedding(32, 100, 16)
    pc = syn.make_field_params(41, "coarse", weight_scale=scale)
    pf = syn.make_field_params(42, "fine", True, True, weight_scale=scale)
    named = [("coarse." + k, v) for k, v in pc.items()] + [("fine." + k, v) for k, v in pf.items()] + [("emb_a", ea), ("emb_t", et)]
    for _, v in named:
        v.requires_grad_(True)
    out = O.render_rays(pc, pf, ea, et, rays, ts, n_samples=64, n_importance=64, perturb=1.0, noise_std=0.0, white_back=True,
                        encode_a=True, encode_t=True, beta_min=0.1, rng=rng, bf16_operands=bf)
    loss = sum(O.nerfw_loss(out, tgt).values())
    gs = torch.autograd.grad(loss, [v for _, v in named])
    return loss.item(), {k: g for (k, _), g in zip(named, gs)}


def product(rays, ts, tgt, rng, scale):
    def build(seed, typ, a=False, t=False):
        m = NeRF(typ, encode_appearance=a, encode_transient=t, beta_min=0.1)
        m.load_state_dict(syn.make_field_params(seed, typ, a, t, weight_scale=scale))
        return m.to(DEV)
    models = {"coarse": build(41, "coarse"), "fine": build(42, "fine", True, True)}
    emb = {"xyz": PosEmbedding(9, 10), "dir": PosEmbedding(3, 4),
           "a": torch.nn.Embedding.from_pretrained(syn.make_embedding(31, 100, 48), freeze=False).to(DEV),
           "t": torch.nn.Embedding.from_pretrained(syn.make_embedding(32, 100, 16), freeze=False).to(DEV)}
    out = render_rays(models, emb, rays.to(DEV), ts.to(DEV), 64, False, 1.0, 0.0, 64, 32768, True,
                      rng={k: v.to(DEV) for k, v in rng.items()})
    loss = sum(O.nerfw_loss(out, tgt.to(DEV)).values())
    named = [(f"coarse.{k}", p) for k, p in models["coarse"].named_parameters()] + \
            [(f"fine.{k}", p) for k, p in models["fine"].named_parameters()] + [("emb_a", emb["a"].weight), ("emb_t", emb["t"].weight)]
    gs = torch.autograd.grad(loss, [p for _, p in named])
    return loss.item(), {k: g.cpu() for (k, _), g in zip(named, gs)}


def tot(a, b):
    return math.sqrt(sum(((a[k] - b[k]).double().norm() ** 2).item() for k in b) / sum((b[k].double().norm() ** 2).item() for k in b))


for n, scale in ((48, 1.0), (512, 1.0), (48, 3.0)):
    rays, ts, tgt, rng = inputs(n)
    lp, gp = product(rays, ts, tgt, rng, scale)
    lf, gf = oracle(rays, ts, tgt, rng, scale, False)
    lb, gb = oracle(rays, ts, tgt, rng, scale, True)
    print(f"=== n={n} scale={scale}: loss product {lp:.7f} fp32 {lf:.7f} bf16model {lb:.7f}")
    print(f"whole-gradient rel err: product-vs-fp32 {tot(gp, gf):.5f}  product-vs-model {tot(gp, gb):.5f}  model-vs-fp32 {tot(gb, gf):.5f}")
    for k in gf:
        r = lambda a, b: ((a[k] - b[k]).double().norm() / b[k].double().norm()).item()
        print(f"{k:36s} |g| {gf[k].norm().item():.3e}  p/f32 {r(gp, gf):.4f}  p/model {r(gp, gb):.4f}  model/f32 {r(gb, gf):.4f}")
