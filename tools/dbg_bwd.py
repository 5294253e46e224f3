import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_fl_b200 import kernels as K, synthetic as syn
from oracle import nerfw_oracle as O
DEV="cuda"
import torch.nn.functional as F
class _R(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x): return x.to(torch.bfloat16).float()
    @staticmethod
    def backward(ctx, g): return g
rb16 = _R.apply
def emu_forward(p, xyz, dir_a, tt, transient):
    """bf16-operand / fp32-accumulate emulation of the fused kernel (per-ray parts stay fp32)."""
    lin = lambda n, x: F.linear(rb16(x), rb16(p[n + ".weight"]), p[n + ".bias"])
    X = xyz
    h = X
    for i in range(8):
        if i == 4: h = torch.cat([X, h], 1)
        h = torch.relu(lin(f"xyz_encoding_{i+1}.0", h))
    sigma = F.softplus(F.linear(h, p["static_sigma.0.weight"], p["static_sigma.0.bias"]))
    feat = lin("xyz_encoding_final", h)
    Wd = p["dir_encoding.0.weight"]
    hd = torch.relu(F.linear(rb16(feat), rb16(Wd[:, :256])) + F.linear(dir_a, Wd[:, 256:], p["dir_encoding.0.bias"]))
    rgb = torch.sigmoid(F.linear(hd, p["static_rgb.0.weight"], p["static_rgb.0.bias"]))
    out = [rgb, sigma]
    if transient:
        Wt = p["transient_encoding.0.weight"]
        ht = torch.relu(F.linear(rb16(feat), rb16(Wt[:, :256])) + F.linear(tt, Wt[:, 256:], p["transient_encoding.0.bias"]))
        for j in (2, 4, 6): ht = torch.relu(lin(f"transient_encoding.{j}", ht))
        out += [torch.sigmoid(F.linear(ht, p["transient_rgb.0.weight"], p["transient_rgb.0.bias"])),
                F.softplus(F.linear(ht, p["transient_sigma.0.weight"], p["transient_sigma.0.bias"])),
                F.softplus(F.linear(ht, p["transient_beta.0.weight"], p["transient_beta.0.bias"]))]
    return torch.cat(out, 1)
def run(n_rays,S,transient,scale):
    gen = torch.Generator().manual_seed(n_rays * 7 + S)
    rays = syn.blender_rays(n_rays, seed=3)
    z = torch.sort(torch.rand(n_rays, S, generator=gen) * 4 + 2, dim=-1)[0]
    sd = syn.make_field_params(33, "fine", True, transient, weight_scale=scale)
    n_vocab = 7
    emb_a, emb_t = syn.make_embedding(5, n_vocab, 48), syn.make_embedding(6, n_vocab, 16)
    ts = torch.randint(0, n_vocab, (n_rays,), generator=gen)
    C9 = 9 if transient else 4
    d_raw = torch.randn(n_rays * S, C9, generator=gen)
    sd_o = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    ea_o, et_o = emb_a.clone().requires_grad_(True), emb_t.clone().requires_grad_(True)
    xyz = (rays[:, None, 0:3] + rays[:, None, 3:6] * z[:, :, None]).reshape(-1, 3)
    parts = [O.posenc(xyz, 10), O.posenc(rays[:, 3:6], 4).repeat_interleave(S, 0), ea_o[ts].repeat_interleave(S, 0)]
    if transient: parts.append(et_o[ts].repeat_interleave(S, 0))
    if EMU:
        raw_o = emu_forward(sd_o, parts[0], torch.cat(parts[1:3], -1), parts[3] if transient else None, transient)
    else:
        raw_o = O.field_forward(sd_o, torch.cat(parts, -1), in_dir_a=27 + 48, output_transient=transient)
    (raw_o * d_raw).sum().backward()
    desc = K.field_desc(63, 27, 48, 16 if transient else 0)
    sdd = {k: v.to(DEV).contiguous() for k, v in sd.items()}
    packed = K.field_pack(desc, sdd)
    ts_d, ea_d, et_d = ts.to(DEV), emb_a.to(DEV), emb_t.to(DEV)
    feat, rb, _ = K.ray_consts_fwd(desc, sdd, rays[:, 3:6].to(DEV).contiguous(), n_freq_dir=4, ts=ts_d, emb_a=ea_d, emb_t=et_d if transient else None, transient=transient)
    mode = K.mlp_mode(transient=transient, save=True)
    saved = K.aligned_empty(K.mlp_save_bytes(desc, n_rays * S, mode), DEV, 16)
    raw = K.mlp_fwd(desc, packed, n_rays=n_rays, S=S, mode=mode, ray_bias=rb, rays=rays.to(DEV), z=z.to(DEV), saved=saved)
    grads = {k: torch.zeros_like(v) for k, v in sdd.items()}
    d_rb = torch.zeros(n_rays, 256, device=DEV)
    K.mlp_bwd(desc, packed, raw, d_raw.to(DEV), saved, n_rays=n_rays, S=S, mode=mode, grads=grads, d_ray_bias=d_rb)
    d_ea, d_et = torch.zeros_like(ea_d), torch.zeros_like(et_d)
    K.ray_consts_bwd(desc, sdd, feat, d_rb, ts=ts_d, n_vocab=n_vocab, transient=transient, grads=grads, d_emb_a=d_ea, d_emb_t=d_et if transient else None)
    torch.cuda.synchronize()
    print(f"--- n_rays={n_rays} S={S} transient={transient} scale={scale}")
    for k, v in sd_o.items():
        ref = v.grad; got = grads[k].cpu()
        err = (got - ref).abs().max().item() / max(ref.abs().max().item(), 1e-9)
        cos = torch.nn.functional.cosine_similarity(got.flatten().double(), ref.flatten().double(), dim=0).item()
        print(f"{k:32s} relerr {err:9.4f} cos {cos:8.5f} |ref| {ref.abs().max().item():.3e} ratio {got.norm().item()/max(ref.norm().item(),1e-12):.4f}")
    for name, got, ref in (("emb_a", d_ea, ea_o.grad), ("emb_t", d_et, et_o.grad)):
        if ref is None: continue
        got = got.cpu()
        err = (got - ref).abs().max().item() / max(ref.abs().max().item(), 1e-9)
        print(f"{name:32s} relerr {err:9.4f} ratio {got.norm().item()/max(ref.norm().item(),1e-12):.4f}")
EMU=True
run(6,64,True,1.0)
run(41,128,True,3.0)
