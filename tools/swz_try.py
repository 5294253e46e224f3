"""One descriptor experiment per process (a bad descriptor faults the context): python tools/swz_try.py <variant> <N> <K> [11 overrides]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_fl_b200 import kernels as K, layout
DEV = "cuda"
variant, N, Kd = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
ov = [int(x) for x in sys.argv[4:15]] if len(sys.argv) > 4 else None
gen = torch.Generator().manual_seed(N * 11 + Kd)
X = torch.randn(Kd, 128, generator=gen).to(torch.float8_e5m2)
Y = torch.randn(Kd, N, generator=gen).to(torch.float8_e4m3fn)
img = {3: layout.tile_image8, 4: layout.tile_image8s32}[variant]
a = img(X.float(), torch.float8_e5m2).to(DEV); b = img(Y.float(), torch.float8_e4m3fn).to(DEV)
want = X.float().T @ Y.float()
d = K.selftest_umma(a, b, N, Kd, variant, ov)
torch.cuda.synchronize()
print("variant", variant, "N", N, "K", Kd, "ov", ov, "maxerr", (d.cpu() - want).abs().max().item(), "ref scale", want.abs().max().item(), flush=True)
