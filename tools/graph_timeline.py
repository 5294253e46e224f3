"""Timeline of ONE replay of the captured training step: a one-thread kernel writes %globaltimer before and after every
C-ABI call of the iteration (tools/csrc/stamp.cu, built by this script), the step is captured with those stamps inside,
replayed, and the stamps of the last replay are printed per call with the stream they were captured on.

The stamps cost ~2 us of stream time each, so the replay is slower than the product's (printed next to the unstamped
time); what the table is for is the ORDER and OVERLAP of the branches and where a stream sits idle.
    python tools/graph_timeline.py [batch]        (one GPU)"""
import ctypes as C
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from nerf_fl_b200 import kernels as K, synthetic as syn
from nerf_fl_b200.nerf import NeRF, PosEmbedding
from nerf_fl_b200.objective import NerfWLoss
from nerf_fl_b200.parallel import FlatTrainState
from nerf_fl_b200.train_step import TrainStep

so = os.path.join(ROOT, "tools", "stamp.so")
if not os.path.exists(so):
    subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-Xcompiler", "-fPIC", "-o", so,
                           os.path.join(ROOT, "tools", "csrc", "stamp.cu")])
lib = C.CDLL(so)
lib.nwb_tool_stamp.argtypes = [C.c_void_p, C.c_int, C.c_void_p]

dev = torch.device("cuda:0")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
slots = torch.zeros(4096, dtype=torch.int64, device=dev)
calls = []          # (name, mode, stream id, slot of the stamp before, slot of the stamp after)
recording = False


def stamp():
    i = stamp.next
    stamp.next += 1
    rc = lib.nwb_tool_stamp(slots.data_ptr(), i, torch.cuda.current_stream(dev).cuda_stream)
    assert rc == 0, rc
    return i


stamp.next = 0
enter0, exit0 = K._timed.__enter__, K._timed.__exit__


def enter(self):
    r = enter0(self)
    self._s0 = stamp() if (recording and torch.cuda.is_current_stream_capturing()) else None     # the capture's calls only
    return r


def exit_(self, *exc):
    if self._s0 is not None:
        calls.append((self.name, self.mode, torch.cuda.current_stream(dev).cuda_stream, self._s0, stamp()))
    return exit0(self, *exc)


def build():
    models = {"coarse": NeRF("coarse"), "fine": NeRF("fine", encode_appearance=True, encode_transient=True, beta_min=0.1)}
    models["coarse"].load_state_dict(syn.make_field_params(41, "coarse"))
    models["fine"].load_state_dict(syn.make_field_params(42, "fine", True, True))
    models = {k: m.to(dev) for k, m in models.items()}
    emb = {"xyz": PosEmbedding(9, 10), "dir": PosEmbedding(3, 4),
           "a": torch.nn.Embedding.from_pretrained(syn.make_embedding(31, 100, 48), freeze=False).to(dev),
           "t": torch.nn.Embedding.from_pretrained(syn.make_embedding(32, 100, 16), freeze=False).to(dev)}
    state = FlatTrainState(list(models.values()) + [emb["a"], emb["t"]], lr=5e-4)
    return TrainStep(models, emb, state, NerfWLoss(), batch=n, N_samples=64, N_importance=64, perturb=1.0, noise_std=0.0,
                     white_back=True)


g = torch.Generator().manual_seed(1)
batches = [(syn.blender_rays(n, seed=s).to(dev), torch.randint(0, 100, (n,), generator=g).to(dev),
            torch.rand(n, 3, generator=g).to(dev)) for s in range(1, 9)]


def time_step(step, iters=100):
    for i in range(10):
        step(*batches[i % 8])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        step(*batches[i % 8])
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


plain = time_step(build())
K._timed.__enter__, K._timed.__exit__ = enter, exit_
step = build()
recording = True
stamped = time_step(step)
recording = False
torch.cuda.synchronize()
t = slots.cpu().tolist()
t0 = min(t[c[3]] for c in calls)
streams = {}
for c in calls:
    streams.setdefault(c[2], len(streams))
print(f"batch {n}: graph replay {plain * 1e3:.1f} us per step; with {stamp.next} stamps inside {stamped * 1e3:.1f} us")
print(f"{'call':28s} {'mode':>5s} {'stream':>6s} {'start us':>9s} {'end us':>9s} {'dur us':>8s}")
for name, mode, s, a, b in sorted(calls, key=lambda c: t[c[3]]):
    print(f"{name:28s} {mode:5d} {streams[s]:6d} {(t[a] - t0) / 1e3:9.1f} {(t[b] - t0) / 1e3:9.1f} {(t[b] - t[a]) / 1e3:8.1f}")
end = max(t[c[4]] for c in calls)
print(f"first stamp -> last stamp: {(end - t0) / 1e3:.1f} us")
