"""Micro-benchmark of the fused field forward (CUDA events, L2-flush between iterations)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_fl_b200 import kernels as K, synthetic as syn

DEV = "cuda"
def bench(n_rays, S, transient, save, iters=10):
    rays = syn.blender_rays(n_rays, seed=3).to(DEV)
    z = (torch.rand(n_rays, S, device=DEV) * 4 + 2).sort(-1)[0]
    sd = {k: v.to(DEV) for k, v in syn.make_field_params(31, "fine", True, transient).items()}
    desc = K.field_desc(63, 27, 48, 16 if transient else 0)
    packed = K.field_pack(desc, sd)
    rb = torch.randn(n_rays, 256, device=DEV)
    mode = K.mlp_mode(transient=transient, save=save)
    saved = K.aligned_empty(K.mlp_save_bytes(desc, n_rays * S, mode), DEV, 16) if save else None
    out = torch.empty(n_rays * S, 9 if transient else 4, device=DEV)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    ts = []
    for i in range(iters + 3):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        K.mlp_fwd(desc, packed, n_rays=n_rays, S=S, mode=mode, ray_bias=rb, rays=rays, z=z, saved=saved, out=out)
        e1.record(); torch.cuda.synchronize()
        if i >= 3: ts.append(e0.elapsed_time(e1))
    ms = sorted(ts)[len(ts) // 2]
    macs = 684160 if transient else 599552
    tf = 2 * macs * n_rays * S / (ms * 1e-3) / 1e12
    print(f"rays={n_rays} S={S} transient={transient} save={save}: {ms:.3f} ms  {tf:.1f} TFLOP/s  {n_rays*S/ms/1e3:.2f} Mpts/s")

if __name__ == "__main__":
    if len(sys.argv) > 1:          # python tools/perf_fwd.py <n_rays> <save 0|1> [iters]   (profiling runs)
        bench(int(sys.argv[1]), 128, True, bool(int(sys.argv[2])), iters=int(sys.argv[3]) if len(sys.argv) > 3 else 2)
    else:
        for n, S in [(1024, 128), (8192, 128), (32768, 128)]:
            for save in (False, True):
                bench(n, S, True, save)
