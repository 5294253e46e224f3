"""Micro-benchmark of the HBM-side kernels at a size where they are not launch-bound (16 384 rays, 128+128 samples):
CUDA events, L2 flushed between iterations, algorithmic bytes of SURVEY.md 8(d) / DESIGN.md section 4 over the median time,
against the measured copy bandwidth in MEASURED_PEAKS.json.      python tools/perf_hbm.py [n_rays]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_fl_b200 import kernels as K, synthetic as syn
DEV = "cuda"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    PEAK = 6500.0
flush = torch.zeros(256 << 20, dtype=torch.uint8, device=DEV)
flush_sink = torch.zeros((), dtype=torch.int64, device=DEV)
ITERS = int(os.environ.get("NWB_PERF_ITERS", "10"))      # 1 under ncu (one profiled launch per kernel and call site)
WARM = 3 if ITERS > 1 else 0
def timed(fn, iters=ITERS):
    ts = []
    for i in range(iters + WARM):
        # L2 flushed by streaming 256 MB THROUGH it with loads: flush.zero_() would leave the cache full of dirty lines
        # whose write-back then competes with the timed kernel's own traffic (measured: +8 us on a 38 us launch)
        flush_sink.copy_(flush.view(torch.int64).sum())
        # the device spins ~150 us here so that the host has enqueued e0, the kernel and e1 before the first of them runs:
        # the interval is the kernel plus the events' own ~2 us, not the host's launch latency
        torch.cuda._sleep(300000)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        if i >= WARM: ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]
def line(name, ms, nbytes):
    gbs = nbytes / (ms * 1e-3) / 1e9
    print(f"{name:34s} {ms * 1e3:8.1f} us  {gbs:7.0f} GB/s  {gbs / PEAK:5.2f} of the measured copy peak ({PEAK:.0f} GB/s)", flush=True)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
Ns = Ni = 128
g = torch.Generator().manual_seed(1)
rays = syn.blender_rays(n, seed=3).to(DEV)
line("sample_coarse (Philox jitter)", timed(lambda: K.sample_coarse(rays, Ns, False, 1.0, None, 7)), n * (32 + 4 * Ns))
z = K.sample_coarse(rays, Ns, False, 1.0, None, 7)
w = (torch.rand(n, Ns, generator=g) ** 4).to(DEV)
line("sample_pdf_merge (stochastic)", timed(lambda: K.sample_pdf_merge(z, w, Ni, False, None, 9)), n * 4 * (2 * Ns + Ns + Ni))
line("sample_pdf_merge (det)", timed(lambda: K.sample_pdf_merge(z, w, Ni, True, None, 9)), n * 4 * (2 * Ns + Ns + Ni))
for C, S, typ, tr in ((4, Ns, "coarse", False), (9, Ns + Ni, "fine", True)):
    raw = torch.rand(n, S, C, generator=g).to(DEV)
    zz = torch.sort(torch.rand(n, S, generator=g) * 4 + 2, dim=-1)[0].to(DEV)
    fb = (24 * S + 20) if C == 4 else (44 * S + 48)
    bb = (36 * S + 32) if C == 4 else (76 * S + 48)
    line(f"composite_fwd<{C}> S={S}", timed(lambda: K.composite_fwd(raw, zz, typ=typ, transient=tr, white_back=True, test_time=False, beta_min=0.1)), n * fb)
    res = K.composite_fwd(raw, zz, typ=typ, transient=tr, white_back=True, test_time=False, beta_min=0.1)
    grads = {k: torch.randn_like(v) for k, v in res.items() if k.startswith(("rgb", "_rgb", "beta", "transient_sigmas", "opacity"))}
    if not tr:
        grads = {k: v for k, v in grads.items() if k in (f"rgb_{typ}",)}
    line(f"composite_bwd<{C}> S={S}", timed(lambda: K.composite_bwd(raw, zz, grads, typ=typ, transient=tr, white_back=True)), n * bb)
    for gen, what in ((1, "ring kernel"), (2, "whole ray resident")):
        K.composite_bwd_select(gen)
        line(f"composite_bwd<{C}> S={S} ({what})", timed(lambda: K.composite_bwd(raw, zz, grads, typ=typ, transient=tr, white_back=True)), n * bb)
    K.composite_bwd_select(0)
