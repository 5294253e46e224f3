"""The optimizer launch (nwb_adam_pack_step_dev: Adam on the flat buffers + refresh of the bf16 operand images) timed alone:
CUDA events, L2 flushed with loads, device spin ahead of the events.     python tools/perf_adam.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_fl_b200 import synthetic as syn
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
state.flat_g.normal_()
flush = torch.zeros(256 << 20, dtype=torch.uint8, device=dev)
sink = torch.zeros((), dtype=torch.int64, device=dev)
ts = []
for i in range(25):
    sink.copy_(flush.view(torch.int64).sum())
    torch.cuda._sleep(300000)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); state.step(); e1.record(); torch.cuda.synchronize()
    if i >= 5: ts.append(e0.elapsed_time(e1))
ts.sort()
print(f"{os.environ.get('NWB_LIB_PATH', 'default'):40s} optimizer step over {state.flat_p.numel()} parameters: median {ts[len(ts) // 2] * 1e3:.1f} us, min {ts[0] * 1e3:.1f} us")
# the plain launch (nwb_adam_step_dev: the same update without the operand-image refresh) and a copy of the same traffic
b = FlatTrainState([m for m in models.values()] + emb, lr=5e-4, fused_pack=False)
def timed(fn):
    ts = []
    for i in range(25):
        sink.copy_(flush.view(torch.int64).sum())
        torch.cuda._sleep(300000)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        if i >= 5: ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2] * 1e3
print(f"plain nwb_adam_step_dev: {timed(b.adam_step):.1f} us")
x = torch.empty(4 * state.flat_p.numel(), device=dev); y = torch.empty_like(x)
print(f"torch copy of {x.numel() * 4 / 1e6:.1f} MB: {timed(lambda: y.copy_(x)):.1f} us;  empty launch: {timed(lambda: torch.cuda._sleep(1)):.1f} us")
