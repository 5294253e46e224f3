"""Golden vectors for the phototourism-notebook shape (test_phototourism.ipynb cell 1: N_emb_xyz = 15 ->
in_channels_xyz = 93, N_vocab 1500, N_a 48, N_tau 16, beta_min 0.03, 128+128 samples), produced by the UNMODIFIED
reference (read-only import from /root/reference).  Build container only:

    python oracle/make_golden_xyz15.py        -> tests/golden/xyz15.npz

Kept apart from make_golden.py so that the files it wrote stay byte-identical."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

from models.nerf import NeRF, PosEmbedding            # noqa: E402  (reference)
from models.rendering import render_rays              # noqa: E402  (reference)
from losses import NerfWLoss                          # noqa: E402  (reference)

from nerf_fl_b200 import synthetic as syn              # noqa: E402

IX = 93


def ref_field(seed, typ, encode_a, encode_t, beta_min=0.03):
    m = NeRF(typ, in_channels_xyz=IX, encode_appearance=encode_a, encode_transient=encode_t, beta_min=beta_min)
    m.load_state_dict(syn.make_field_params(seed, typ, encode_a, encode_t, in_xyz=IX))
    return m


def run_ref(models, emb, rays, ts, rngt, **kw):
    queue = list(rngt)
    rl, rnl, r = torch.rand_like, torch.randn_like, torch.rand
    torch.rand_like = lambda t, **k: queue.pop(0)
    torch.randn_like = lambda t, **k: queue.pop(0)
    torch.rand = lambda *a, **k: queue.pop(0)
    try:
        return render_rays(models, emb, rays, ts, **kw)
    finally:
        torch.rand_like, torch.randn_like, torch.rand = rl, rnl, r


arrs = {}
# ---- NeRF.forward on explicit encoded rows, four head modes (models/nerf.py:153-212 with in_channels_xyz = 93) ----
g = torch.Generator().manual_seed(112)
xin = torch.rand(200, IX + 27 + 48 + 16, generator=g) * 2 - 1
arrs["field/x"] = xin
arrs["field/sigma_only"] = ref_field(21, "coarse", False, False)(xin[:, :IX], sigma_only=True)
arrs["field/plain"] = ref_field(21, "coarse", False, False)(xin[:, :IX + 27], output_transient=False)
arrs["field/a"] = ref_field(22, "fine", True, False)(xin[:, :IX + 27 + 48], output_transient=False)
arrs["field/at"] = ref_field(23, "fine", True, True)(xin, output_transient=True)

# ---- render_rays end to end: the notebook's test_time call and a training step with gradients ----------------------
N, Ns, Ni, NV = 40, 128, 128, 1500
rays = syn.blender_rays(N, seed=7)
g = torch.Generator().manual_seed(115)
ts = torch.randint(0, NV, (N,), generator=g)
rng_t = {"perturb_rand": torch.rand(N, Ns, generator=g), "noise_coarse": torch.randn(N, Ns, generator=g),
         "u": torch.rand(N, Ni, generator=g)}
emb_a = syn.make_embedding(31, NV, 48)
emb_t = syn.make_embedding(32, NV, 16)
emb = {"xyz": PosEmbedding(14, 15), "dir": PosEmbedding(3, 4),
       "a": torch.nn.Embedding.from_pretrained(emb_a, freeze=False),
       "t": torch.nn.Embedding.from_pretrained(emb_t, freeze=False)}
arrs.update({"rays": rays, "ts": ts, **{"rng_" + k: v for k, v in rng_t.items()}})
models = {"coarse": ref_field(41, "coarse", False, False), "fine": ref_field(42, "fine", True, True, beta_min=0.03)}
with torch.no_grad():
    out = run_ref(models, emb, rays, ts, [rng_t["noise_coarse"]], N_samples=Ns, use_disp=False, perturb=0, noise_std=0,
                  N_importance=Ni, chunk=32768, white_back=False, test_time=True)
for k, v in out.items():
    arrs[f"test/{k}"] = v
out = run_ref(models, emb, rays, ts, [rng_t["perturb_rand"], rng_t["noise_coarse"], rng_t["u"]], N_samples=Ns, use_disp=False,
              perturb=1.0, noise_std=0.0, N_importance=Ni, chunk=32768, white_back=False, test_time=False)
for k, v in out.items():
    arrs[f"train/{k}"] = v
tgt = torch.rand(N, 3, generator=torch.Generator().manual_seed(3))
arrs["targets"] = tgt
loss = sum(NerfWLoss()(out, tgt).values())
params = [(f"coarse.{k}", p) for k, p in models["coarse"].named_parameters()] + \
         [(f"fine.{k}", p) for k, p in models["fine"].named_parameters()] + \
         [("emb_a", emb["a"].weight), ("emb_t", emb["t"].weight)]
grads = torch.autograd.grad(loss, [p for _, p in params])
arrs["train/loss"] = loss.detach()
for (k, _), gr in zip(params, grads):
    arrs[f"train/grad_norm/{k}"] = gr.norm()
    if gr.numel() <= 32768 and not k.startswith("emb"):
        arrs[f"train/grad/{k}"] = gr.clone()           # whole small tensors (first layers: the 93-column operands)

out_path = os.path.join(ROOT, "tests", "golden", "xyz15.npz")
np.savez_compressed(out_path, **{k: (v.detach().numpy() if torch.is_tensor(v) else np.asarray(v)) for k, v in arrs.items()})
print("wrote", out_path, len(arrs), "arrays", os.path.getsize(out_path), "bytes")
