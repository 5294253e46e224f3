"""Timeline of block 0 / tile 0 of the fused field forward (clock64 stamps written by the kernel's debug hook)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_fl_b200 import kernels as K, synthetic as syn, _lib as L

DEV = "cuda"
n_rays, S, transient = 1024, 128, True
save = len(sys.argv) > 1 and sys.argv[1] == "1"
rays = syn.blender_rays(n_rays, seed=3).to(DEV)
z = (torch.rand(n_rays, S, device=DEV) * 4 + 2).sort(-1)[0]
sd = {k: v.to(DEV) for k, v in syn.make_field_params(31, "fine", True, transient).items()}
desc = K.field_desc(63, 27, 48, 16)
packed = K.field_pack(desc, sd)
rb = torch.randn(n_rays, 256, device=DEV)
mode = K.mlp_mode(transient=transient, save=save)
saved = K.aligned_empty(K.mlp_save_bytes(desc, n_rays * S, mode), DEV, 16) if save else None
tr = torch.zeros(16 * 13, dtype=torch.int64, device=DEV)
fn = L.lib().nwb_debug_set_fwd_trace
fn.argtypes = [C.c_void_p]; fn.restype = None
for i in range(3):
    if i == 2:
        fn(tr.data_ptr())
    K.mlp_fwd(desc, packed, n_rays=n_rays, S=S, mode=mode, ray_bias=rb, rays=rays, z=z, saved=saved)
torch.cuda.synchronize()
fn(None)
t = tr.cpu().view(13, 16)
t0 = int(t[0, 0])
print("cycles relative to layer 0 operand-ready; per layer: MMA ready@kb0..4 | issued@kb0..4 | epi acc seen | quarters handed 0..3")
for g in range(13):
    r = [int(x) - t0 if int(x) else None for x in t[g]]
    print(f"g={g:2d} ready {r[0:5]} issued {r[5:10]} acc {r[10]} q {r[11:15]}")
