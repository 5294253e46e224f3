"""Per-tensor gradient errors of NeRF.forward's backward (random cotangents) vs the oracle's bf16-operand model."""
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
sdb = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
(O.field_forward_bf16(sdb, x[:, :63], x[:, 63:138], x[:, 138:], True) * d).sum().backward()
for k, p in m.named_parameters():
    got, want = p.grad.double().cpu(), sdb[k].grad.double()
    print(f"{k:36s} |g| {want.norm().item():.3e}  rel {((got - want).norm() / want.norm()).item():.4f}")
