"""Micro-benchmark of the three field kernels (forward saving / not saving, dgrad, wgrad) at the shapes of a 1024-ray
training step and of an inference chunk: CUDA events, L2 flushed between iterations, median of `iters`.
    python tools/perf_field.py [iters]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_fl_b200 import kernels as K, synthetic as syn

DEV = "cuda"
flush = None

def timed(fn, iters):
    global flush
    if flush is None:
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    ts = []
    for i in range(iters + 3):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        if i >= 3: ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]

def bench(n_rays, S, fine, iters, what=("fwd", "fwd_save", "dgrad", "wgrad")):
    transient = fine
    rays = syn.blender_rays(n_rays, seed=3).to(DEV)
    z = (torch.rand(n_rays, S, device=DEV) * 4 + 2).sort(-1)[0]
    sd = {k: v.to(DEV) for k, v in syn.make_field_params(31, "fine" if fine else "coarse", fine, transient).items()}
    desc = K.field_desc(63, 27, 48 if fine else 0, 16 if transient else 0)
    packed = K.field_pack(desc, sd)
    rb = torch.randn(n_rays, 256, device=DEV)
    C = 9 if transient else 4
    out = torch.empty(n_rays * S, C, device=DEV)
    macs = 684160 if transient else 593408
    res = {}
    mode_s = K.mlp_mode(transient=transient, save=True)
    need_save = any(w in what for w in ("fwd_save", "dgrad", "wgrad"))
    saved = K.aligned_empty(K.mlp_save_bytes(desc, n_rays * S, mode_s), DEV, 16) if need_save else None
    if "fwd" in what:
        mode = K.mlp_mode(transient=transient)
        res["fwd"] = timed(lambda: K.mlp_fwd(desc, packed, n_rays=n_rays, S=S, mode=mode, ray_bias=rb, rays=rays, z=z, out=out), iters)
    if "fwd_save" in what or "dgrad" in what or "wgrad" in what:
        res["fwd_save"] = timed(lambda: K.mlp_fwd(desc, packed, n_rays=n_rays, S=S, mode=mode_s, ray_bias=rb, rays=rays, z=z, saved=saved, out=out), iters)
    if "dgrad" in what or "wgrad" in what:
        d_raw = torch.randn_like(out) * 1e-3
        grads = {k: torch.zeros_like(v) for k, v in sd.items()}
        drb = torch.zeros(n_rays, 256, device=DEV)
        ws = K.aligned_empty(K.mlp_bwd_workspace_bytes(desc, n_rays * S, mode_s), DEV, 16)
        res["dgrad"] = timed(lambda: K.mlp_bwd(desc, packed, out, d_raw, saved, n_rays=n_rays, S=S, mode=mode_s, grads=grads, d_ray_bias=drb, workspace=ws, phases=("dgrad",)), iters)
        res["wgrad"] = timed(lambda: K.mlp_bwd(desc, packed, out, d_raw, saved, n_rays=n_rays, S=S, mode=mode_s, grads=grads, d_ray_bias=drb, workspace=ws, phases=("wgrad",)), iters)
    line = f"rays={n_rays} S={S} {'fine' if fine else 'coarse'}:"
    for k, ms in res.items():
        fl = 2 * macs * n_rays * S * (1.0 if k != "dgrad" else (macs - 35712) / macs)
        line += f"  {k} {ms:.4f} ms ({fl / (ms * 1e-3) / 1e12:.0f} TF)"
    print(line, flush=True)

if __name__ == "__main__":
    iters = int(sys.argv[1]) if len(sys.argv) > 1 else 10
    print("lib:", os.environ.get("NWB_LIB_PATH", "default"))
    bench(1024, 64, False, iters)
    bench(1024, 128, True, iters)
    bench(32768, 128, True, iters, what=("fwd",))
