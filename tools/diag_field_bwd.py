"""Per-tensor gradient errors of NeRF.forward's backward (random cotangents) vs the oracle's operand-rounding models:
q8 (bf16 GEMM operands + 8-bit weight-gradient images, what the kernels do), bf16-only, and plain fp32."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_fl_b200 import synthetic as syn
from nerf_fl_b200.nerf import NeRF
from oracle import nerfw_oracle as O

n_tiles = int(sys.argv[1]) if len(sys.argv) > 1 else 64
coherent = len(sys.argv) > 2 and sys.argv[2] == "coherent"
n_points = n_tiles * 128 - 5
g = torch.Generator().manual_seed(n_tiles)
x = torch.rand(n_points, 154, generator=g) * 2 - 1
d = torch.randn(n_points, 9, generator=g) / n_points
if coherent:
    d = d.abs()
m = NeRF("fine", encode_appearance=True, encode_transient=True)
sd = syn.make_field_params(23, "fine", True, True)
m.load_state_dict(sd)
m.to("cuda")
out = m(x.cuda())
out.backward(d.cuda())
def model(fn):
    sdb = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    if fn is None:
        (O.field_forward(sdb, x, in_dir_a=75) * d).sum().backward()
    else:
        (fn(sdb, x[:, :63], x[:, 63:138], x[:, 138:], True) * d).sum().backward()
    return {k: v.grad.double() for k, v in sdb.items()}


q8, bf, f32 = model(O.field_forward_q8), model(O.field_forward_bf16), model(None)
rel = lambda a, b: ((a - b).norm() / b.norm()).item()
for k, p in m.named_parameters():
    got = p.grad.double().cpu()
    print(f"{k:36s} |g| {q8[k].norm().item():.3e}  vs q8 {rel(got, q8[k]):.4f}  vs bf16 {rel(got, bf[k]):.4f}  vs fp32 {rel(got, f32[k]):.4f}"
          f"   (q8 vs bf16 {rel(q8[k], bf[k]):.4f}, bf16 vs fp32 {rel(bf[k], f32[k]):.4f})")
