"""Per-block timeline of the optimizer launch (a -DNWB_ADAM_TRACE build): when each block starts, passes its prologue and
ends, by grid slice.    python tools/build_variant.py adam_trace -DNWB_ADAM_TRACE; NWB_LIB_PATH=tools/variants/adam_trace.so python tools/trace_adam.py"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_fl_b200 import synthetic as syn, _lib as L
from nerf_fl_b200.nerf import NeRF
from nerf_fl_b200.parallel import FlatTrainState
dev = torch.device("cuda:0")
models = {"coarse": NeRF("coarse"), "fine": NeRF("fine", encode_appearance=True, encode_transient=True, beta_min=0.1)}
models["coarse"].load_state_dict(syn.make_field_params(41, "coarse"))
models["fine"].load_state_dict(syn.make_field_params(42, "fine", True, True))
models = {k: m.to(dev) for k, m in models.items()}
emb = [torch.nn.Embedding.from_pretrained(syn.make_embedding(31, 100, 48), freeze=False).to(dev),
       torch.nn.Embedding.from_pretrained(syn.make_embedding(32, 100, 16), freeze=False).to(dev)]
state = FlatTrainState(list(models.values()) + emb, lr=5e-4)
flush = torch.zeros(256 << 20, dtype=torch.uint8, device=dev)
for i in range(5):
    state.flat_g.normal_()
    flush.view(torch.int64).sum()
    state.step()
torch.cuda.synchronize()
buf = (C.c_ulonglong * (3 * 4096))()
fn = L.lib().nwb_debug_adam_trace
fn.argtypes = [C.c_void_p, C.c_int]
L.check(fn(buf, 3 * 4096), "trace")
t = list(buf)
nb = max(b for b in range(4096) if t[3 * b + 2]) + 1           # blocks of the last launch (24 per grid slice)
assert nb % 24 == 0, nb
t0 = min(t[0:3 * nb:3])
print(f"{nb} blocks; kernel span {(max(t[2:3 * nb:3]) - t0) / 1e3:.1f} us")
print("slice  first start  last start   mean prologue us   mean body us   last end")
for y in range(nb // 24):
    bl = [(t[3 * b] - t0, t[3 * b + 1] - t[3 * b], t[3 * b + 2] - t[3 * b + 1], t[3 * b + 2] - t0) for b in range(y * 24, y * 24 + 24)]
    print(f"{y:5d} {min(b[0] for b in bl) / 1e3:11.1f} {max(b[0] for b in bl) / 1e3:11.1f} {sum(b[1] for b in bl) / 24e3:15.2f} "
          f"{sum(b[2] for b in bl) / 24e3:14.2f} {max(b[3] for b in bl) / 1e3:10.1f}")
