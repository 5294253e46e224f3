"""Timeline of the pair forward kernel (library built with -DNWB_FWD_TRACE): cluster 0 / rank 0, second quad of tiles.
    NWB_LIB_PATH=tools/variants/fwdtrace.so python tools/trace_fwd_dual.py [save 0|1]"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_fl_b200 import kernels as K, synthetic as syn, _lib as L

DEV = "cuda"
n_rays, S, transient = 2048, 128, True
save = len(sys.argv) > 1 and sys.argv[1] == "1"
rays = syn.blender_rays(n_rays, seed=3).to(DEV)
z = (torch.rand(n_rays, S, device=DEV) * 4 + 2).sort(-1)[0]
sd = {k: v.to(DEV) for k, v in syn.make_field_params(31, "fine", True, transient).items()}
desc = K.field_desc(63, 27, 48, 16)
packed = K.field_pack(desc, sd)
rb = torch.randn(n_rays, 256, device=DEV)
mode = K.mlp_mode(transient=transient, save=save)
saved = K.aligned_empty(K.mlp_save_bytes(desc, n_rays * S, mode), DEV, 16) if save else None
tr = torch.zeros(13 * 2 * 8, dtype=torch.int64, device=DEV)
fn = L.lib().nwb_debug_set_fwd_trace
fn.argtypes = [C.c_void_p]; fn.restype = None
K.mlp_fwd_select(2)
for i in range(3):
    if i == 2:
        fn(tr.data_ptr())
    K.mlp_fwd(desc, packed, n_rays=n_rays, S=S, mode=mode, ray_bias=rb, rays=rays, z=z, saved=saved)
torch.cuda.synchronize()
fn(None)
t = tr.cpu().view(13, 2, 8)
t0 = int(t[0, 0, 0])
print(f"pair kernel, save={int(save)}: cycles relative to the MMA thread seeing layer 0 / slot 0")
print(" g t | MMA sees operand | MMAs issued | epi sees acc | slices written | handoff out | stores/heads done")
for g in range(13):
    for s in range(2):
        r = [int(x) - t0 if int(x) else None for x in t[g, s, :6]]
        print(f"{g:2d} {s} | {r[0]} | {r[1]} | {r[2]} | {r[3]} | {r[4]} | {r[5]}")
