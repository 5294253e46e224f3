"""Forward dispatch A/B at a given tile count: auto (pair kernel + one-tile tail) vs pair kernel only vs one-tile only."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_fl_b200 import kernels as K, synthetic as syn
DEV = "cuda"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
def run(n_rays, S, save, gen, iters=10):
    rays = syn.blender_rays(n_rays, seed=3).to(DEV)
    z = (torch.rand(n_rays, S, device=DEV) * 4 + 2).sort(-1)[0]
    sd = {k: v.to(DEV) for k, v in syn.make_field_params(31, "fine", True, True).items()}
    desc = K.field_desc(63, 27, 48, 16)
    packed = K.field_pack(desc, sd)
    rb = torch.randn(n_rays, 256, device=DEV)
    mode = K.mlp_mode(transient=True, save=save)
    saved = K.aligned_empty(K.mlp_save_bytes(desc, n_rays * S, mode), DEV, 16) if save else None
    out = torch.empty(n_rays * S, 9, device=DEV)
    K.mlp_fwd_select(gen)
    ts = []
    for i in range(iters + 3):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); K.mlp_fwd(desc, packed, n_rays=n_rays, S=S, mode=mode, ray_bias=rb, rays=rays, z=z, saved=saved, out=out); e1.record()
        torch.cuda.synchronize()
        if i >= 3: ts.append(e0.elapsed_time(e1))
    K.mlp_fwd_select(0)
    return sorted(ts)[len(ts) // 2]
for n_rays, S in [(1024, 128), (1024, 64), (512, 128), (300, 128), (2048, 128)]:
    for save in (False, True):
        print(f"tiles={n_rays * S // 128:5d} save={int(save)}: " + "  ".join(f"gen{g} {run(n_rays, S, save, g):.4f} ms" for g in (0, 2, 1)), flush=True)
