"""Timing of the field backward kernels (dgrad / wgrad, CUDA events, L2 flushed) and, when the library is built with
-DNWB_BWD_TRACE, the timeline of block 0's first tile in the dgrad kernel (clock64 stamps)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_fl_b200 import kernels as K, synthetic as syn, _lib as L

DEV = "cuda"
n_rays = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
S, transient = 128, True
rays = syn.blender_rays(n_rays, seed=3).to(DEV)
z = (torch.rand(n_rays, S, device=DEV) * 4 + 2).sort(-1)[0]
sd = {k: v.to(DEV) for k, v in syn.make_field_params(31, "fine", True, transient).items()}
desc = K.field_desc(63, 27, 48, 16)
packed = K.field_pack(desc, sd)
rb = torch.randn(n_rays, 256, device=DEV)
mode = K.mlp_mode(transient=transient, save=True)
saved = K.aligned_empty(K.mlp_save_bytes(desc, n_rays * S, mode), DEV, 16)
raw = K.mlp_fwd(desc, packed, n_rays=n_rays, S=S, mode=mode, ray_bias=rb, rays=rays, z=z, saved=saved)
d_raw = torch.randn_like(raw)
grads = {k: torch.zeros_like(v) for k, v in sd.items()}
d_rb = torch.zeros(n_rays, 256, device=DEV)
ws = K.aligned_empty(K.mlp_bwd_workspace_bytes(desc, n_rays * S, mode), DEV, 16)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
tr = torch.zeros(8 * 13 * 2, dtype=torch.int64, device=DEV)
GEN = int(os.environ.get("NWB_DGRAD_GEN", "2"))
K.mlp_dgrad_select(GEN)
fn = L.lib().nwb_debug_set_bwd_trace
fn.argtypes = [C.c_void_p]; fn.restype = None
for ph in ("dgrad", "wgrad"):
    ts = []
    for i in range(8):
        flush.zero_()
        if ph == "dgrad" and i == 7:
            fn(tr.data_ptr())
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        K.mlp_bwd(desc, packed, raw, d_raw, saved, n_rays=n_rays, S=S, mode=mode, grads=grads, d_ray_bias=d_rb,
                  workspace=ws, phases=(ph,))
        e1.record(); torch.cuda.synchronize()
        fn(None)
        if i >= 3: ts.append(e0.elapsed_time(e1))
    ms = sorted(ts)[len(ts) // 2]
    fl = (1296896 if ph == "dgrad" else 1368320) * n_rays * S
    print(f"{ph}: rays={n_rays} S={S}: {ms:.3f} ms  {fl / ms / 1e9:.1f} TFLOP/s  {ms * 1e3 / (n_rays * S / 1e3):.3f} us/kpt")
if GEN == 2 and int(tr.abs().sum()) != 0:
    t = tr.cpu().view(13, 2, 8)
    t0 = int(t[12, 0, 2])
    print("two-slot dgrad kernel, cluster 0 / rank 0, SECOND quad (steady state), cycles from slot 0's layer-12 epilogue start:")
    print(" g t | step top | before acc wait | acc seen | slices written | handoff out | MMA sees dZ | MMAs issued")
    for g in range(12, -1, -1):
        for sl in range(2):
            r = [int(x) - t0 if int(x) else None for x in t[g, sl]]
            print(f"{g:2d} {sl} | {r[6]} | {r[5]} | {r[2]} | {r[3]} | {r[4]} | {r[0]} | {r[1]}")
    sys.exit(0)
t = tr.cpu()[:104].view(13, 8)
if int(t.abs().sum()) != 0:
    t0 = int(t[12, 2])
    print("dgrad block 0 SECOND tile pair (steady state: weights in L2), cycles from the start of layer 12's epilogue:")
    print(" g | acc seen | sD free | leader slices done | all slices done | MMA sees dZ | MMAs issued")
    for g in range(12, -1, -1):
        r = [int(x) - t0 if int(x) else None for x in t[g]]
        print(f"{g:2d} | {r[2]} | {r[3]} | {r[4]} | {r[5]} | {r[0]} | {r[1]}")
