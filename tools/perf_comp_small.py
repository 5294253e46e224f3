"""The compositing backward at the training step's shapes (1024 rays), alone: ring kernel vs whole-ray-resident kernel."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_fl_b200 import kernels as K
DEV = "cuda"
g = torch.Generator().manual_seed(1)
flush = torch.zeros(256 << 20, dtype=torch.uint8, device=DEV); sink = torch.zeros((), dtype=torch.int64, device=DEV)
def timed(fn, iters=15):
    ts = []
    for i in range(iters + 3):
        sink.copy_(flush.view(torch.int64).sum()); torch.cuda._sleep(300000)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        if i >= 3: ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2] * 1e3
n = 1024
print(f"empty launch: {timed(lambda: torch.cuda._sleep(1)):.1f} us")
for C, S, typ, tr in ((4, 64, "coarse", False), (9, 128, "fine", True)):
    raw = torch.rand(n, S, C, generator=g).to(DEV)
    zz = torch.sort(torch.rand(n, S, generator=g) * 4 + 2, dim=-1)[0].to(DEV)
    rgb, tgt = torch.rand(n, 3, generator=g).to(DEV), torch.rand(n, 3, generator=g).to(DEV)
    beta = (torch.rand(n, generator=g) + 0.1).to(DEV)
    one = torch.ones((), device=DEV)
    seed = dict(rgb=rgb, beta=beta if tr else None, targets=tgt, g_terms=[None, None, None, None, one], coef=1.0, lambda_u=0.01,
                n_rays=n, n_ts=n * S if tr else 0, fine=tr)
    amax = torch.zeros(64, dtype=torch.int32, device=DEV)
    for gen, what in ((1, "ring"), (2, "resident")):
        K.composite_bwd_select(gen)
        for sd, am, tag in ((None, None, "plain"), (seed, None, "seeded"), (seed, amax, "seeded+amax")):
            t = timed(lambda: K.composite_bwd(raw, zz, {} if sd else {f"rgb_{typ}": rgb}, typ=typ, transient=tr, white_back=True, seed=sd, amax_out=am))
            print(f"C={C} S={S} {what:9s} {tag:12s} {t:6.1f} us")
K.composite_bwd_select(0)
