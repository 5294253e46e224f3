#!/usr/bin/env python
"""bench.py -- NeRF-W rays/s on B200 (BASELINE.json metric), one JSON line on stdout.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload train|infer]

Default workload = BASELINE.json configs[1]: a NeRF-W training step (render_rays forward, NerfWLoss,
backward, gradient all-reduce when N>1, Adam on the flat buffer, bf16 operand-cache refresh) on
Blender-shaped synthetic rays, batch 1024 per rank, encode_a + encode_t, beta_min 0.1, 64+64 samples.
`value` is whole-job rays/s with the ray batches resident in HBM; `e2e` is the same step driven from
pinned HOST buffers (H2D of rays/ts/targets and D2H of the loss inside the timed region).

`--impl reference` times the reference's CPU arithmetic for the same step on the host cores through
oracle/nerfw_oracle.py (a torch-CPU restatement of models/rendering.py + models/nerf.py + losses.py,
pinned to the reference by tests/golden; the Python reference itself cannot travel to the GPU box).
This and the `cpu_baseline` leg are the only places bench.py touches oracle/.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "NeRF-W rays/sec (train fwd+bwd, infer) 64+64 samples at 1/2/4/8 B200"
N_SAMPLES, N_IMPORTANCE = 64, 64
BATCH = 1024
N_VOCAB = 100
BETA_MIN = 0.1
N_ROTATE = 8          # distinct resident ray batches cycled through (inputs differ every step)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="train", choices=["train", "infer"])
    ap.add_argument("--batch", type=int, default=BATCH, help="rays per rank per step (train)")
    ap.add_argument("--samples", type=int, default=N_SAMPLES, help="coarse samples (= fine importance samples)")
    ap.add_argument("--vocab", type=int, default=N_VOCAB, help="embedding rows (configs[2], Phototourism-shaped: 1500)")
    ap.add_argument("--beta-min", type=float, default=BETA_MIN, help="configs[2]: 0.03")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="launch the training step eagerly instead of replaying its CUDA graph")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------------
# clocks: sample nvidia-smi DURING the timed region (B200_PROFILING.md recipe)
# ---------------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.FIELDS}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        self.thread.join(timeout=2)
        sm, smax, reasons, power = [], [], set(), []
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for ln in self.lines:
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 9:
                continue
            try:
                sm.append(float(p[1]))
                smax.append(float(p[2]))
                power.append(float(p[3]))
            except ValueError:
                continue
            for name, val in zip(names, p[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm), "power_w_max": max(power) if power else None}


# ---------------------------------------------------------------------------------------------------
# workload: synthetic Blender-shaped rays, seeded weights, per-ray ts / targets
# ---------------------------------------------------------------------------------------------------
def host_batches(n_batches: int, batch: int, rank: int):
    import torch
    from nerf_fl_b200 import synthetic as syn
    out = []
    for b in range(n_batches):
        seed = 1000 * rank + b + 1
        gen = torch.Generator().manual_seed(seed)
        rays = syn.blender_rays(batch, seed=seed)
        ts = torch.randint(0, N_VOCAB, (batch,), generator=gen)
        tgt = torch.rand(batch, 3, generator=gen)
        out.append((rays, ts, tgt))
    return out


def train_flops_per_ray(ns: int, ni: int) -> float:
    """training MACs/sample = 2*fwd + (fwd - 35 712) ; coarse-full fwd 593 408,
    fine (a+t) fwd 684 160; FLOP = 2*MAC."""
    coarse, fine = 3 * 593408 - 35712, 3 * 684160 - 35712
    return 2.0 * (ns * coarse + (ns + ni) * fine)


def infer_flops_per_ray(ns: int, ni: int) -> float:
    return 2.0 * (ns * 491264 + (ns + ni) * 684160)


def field_flops_per_point(kind: str, typ: str) -> float:
    """algorithmic FLOPs per sample point of one field kernel launch (no padding counted)."""
    fwd = {"coarse": 593408, "coarse_sigma": 491264, "fine": 684160}[typ]
    if kind == "fwd":
        return 2.0 * fwd
    full = {"coarse": 593408, "fine": 684160}[typ]
    if kind == "dgrad":
        return 2.0 * (full - 35712)
    if kind == "wgrad":
        return 2.0 * full
    raise KeyError(kind)


def build_product(dev, rank: int):
    import torch
    from nerf_fl_b200 import synthetic as syn
    from nerf_fl_b200.nerf import NeRF, PosEmbedding
    models = {"coarse": NeRF("coarse"),
              "fine": NeRF("fine", encode_appearance=True, encode_transient=True, beta_min=BETA_MIN)}
    models["coarse"].load_state_dict(syn.make_field_params(11, "coarse"))
    models["fine"].load_state_dict(syn.make_field_params(12, "fine", True, True))
    for m in models.values():
        m.to(dev)
    emb = {"xyz": PosEmbedding(9, 10), "dir": PosEmbedding(3, 4),
           "a": torch.nn.Embedding.from_pretrained(syn.make_embedding(13, N_VOCAB, 48), freeze=False).to(dev),
           "t": torch.nn.Embedding.from_pretrained(syn.make_embedding(14, N_VOCAB, 16), freeze=False).to(dev)}
    return models, emb


# ---------------------------------------------------------------------------------------------------
# CPU arm: the reference's arithmetic for the same step on the host cores (oracle port, torch CPU)
# ---------------------------------------------------------------------------------------------------
def cpu_train_setup(batch: int):
    import torch
    from nerf_fl_b200 import synthetic as syn
    pc = syn.make_field_params(11, "coarse")
    pf = syn.make_field_params(12, "fine", True, True)
    ea, et = syn.make_embedding(13, N_VOCAB, 48), syn.make_embedding(14, N_VOCAB, 16)
    leaves = list(pc.values()) + list(pf.values()) + [ea, et]
    for v in leaves:
        v.requires_grad_(True)
    opt = torch.optim.Adam(leaves, lr=5e-4, eps=1e-8)          # utils/__init__.py:31
    return pc, pf, ea, et, opt


def cpu_train_step(state, rays, ts, tgt, ns, ni):
    from oracle import nerfw_oracle as O
    pc, pf, ea, et, opt = state
    opt.zero_grad(set_to_none=True)
    out = O.render_rays(pc, pf, ea, et, rays, ts, n_samples=ns, n_importance=ni, perturb=1.0, noise_std=0.0,
                        white_back=True, encode_a=True, encode_t=True, beta_min=BETA_MIN)
    loss = sum(O.nerfw_loss(out, tgt).values())
    loss.backward()
    opt.step()
    return float(loss.detach())


def cpu_infer_step(state, rays, ts, ns, ni):
    import torch
    from oracle import nerfw_oracle as O
    pc, pf, ea, et, _ = state
    with torch.no_grad():
        out = O.render_rays(pc, pf, ea, et, rays, ts, n_samples=ns, n_importance=ni, perturb=0.0, noise_std=0.0,
                            white_back=True, test_time=True, encode_a=True, encode_t=True, beta_min=BETA_MIN)
    return float(out["rgb_fine"].sum())


def cpu_leg(workload: str, ns: int, ni: int, steps: int, warmup: int, budget_s: float, batch: int):
    """Times `steps` CPU steps after `warmup`; the per-step ray count is cut (power of two, >= 64) until the
    whole leg fits `budget_s`.  -> (rays/s, sample description, cores, ms/step, rays per step)."""
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    state = cpu_train_setup(batch)
    rays, ts, tgt = host_batches(1, batch, 0)[0]
    n = batch

    def one(n):
        if workload == "train":
            cpu_train_step(state, rays[:n], ts[:n], tgt[:n], ns, ni)
        else:
            cpu_infer_step(state, rays[:n], ts[:n], ns, ni)

    probe_n = min(n, 128)
    one(probe_n)                                   # page-in / thread-pool warm-up
    t0 = time.perf_counter()
    one(probe_n)
    per_ray = (time.perf_counter() - t0) / probe_n
    while n > 64 and per_ray * n * (steps + warmup) > budget_s:
        n //= 2
    for _ in range(warmup):
        one(n)
    t0 = time.perf_counter()
    for _ in range(steps):
        one(n)
    dt = time.perf_counter() - t0
    what = "training step (render_rays fwd, NerfWLoss, bwd, Adam)" if workload == "train" else "test_time render_rays"
    sample = f"{steps} x {n}-ray {what}, {ns}+{ni} samples, after {warmup} warm-up, torch {torch.__version__} CPU, {cores} threads"
    return n * steps / dt, sample, cores, dt / steps * 1e3, n


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    ns = ni = args.samples
    v, sample, cores, ms, n = cpu_leg(args.workload, ns, ni, args.steps, args.warmup, 150.0, args.batch)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "rays/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, n_rays_step=n),
            "cpu_baseline": {"value": v, "unit": "rays/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


def ncu_traffic(dom, args):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the committed
    `ncu --set full` capture of this same command (profiles/ncu_traffic.json, written by tools/ncu_table.py);
    None when no capture exists for this kernel / workload size."""
    try:
        tab = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
    except (OSError, ValueError):
        return None
    key = f"{dom['kernel']}|mode{dom['mode']}|batch{args.batch}|samples{args.samples}|{args.workload}"
    ent = tab.get(key)
    return None if ent is None else ent["bytes"]          # bytes per launch


def workload_config(args, n_rays_step=None):
    ns = args.samples
    if args.workload == "train":
        name = (f"configs[1]: NeRF-W training step fwd+bwd (+all-reduce, Adam), Blender-shaped synthetic rays 400x400, "
                f"batch {args.batch}/rank, encode_a+encode_t, N_vocab {N_VOCAB}, beta_min {BETA_MIN}, {ns}+{ns} samples, "
                f"perturb 1, noise_std 0, white_back")
    else:
        name = (f"configs[3]-shaped: NeRF-W test_time render_rays, 800x800 synthetic image, chunk 32768, {ns}+{ns} samples, "
                f"ray-sharded across ranks, rgb+depth gathered")
    cfg = {"workload": name, "rays_per_step_per_rank": n_rays_step if n_rays_step is not None else args.batch,
           "n_samples": ns, "n_importance": ns,
           "launch": "eager" if (args.no_graph or args.workload != "train") else "CUDA graph replay of the whole iteration",
           "l2": f"{N_ROTATE} distinct resident ray batches rotate; per-step working set (saved activations + dZ workspace, "
                 f"~2 GB at batch 1024) exceeds the 126 MB L2"}
    return cfg


# ---------------------------------------------------------------------------------------------------
# product arm
# ---------------------------------------------------------------------------------------------------
def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return {"hbm_gbs": p["hbm_gbs"], "bf16_burst": p["bf16_tflops"], "bf16_sustained": p["bf16_tflops_sustained"],
                "source": "MEASURED_PEAKS.json"}
    # B200_PROFILING.md fallback figures
    return {"hbm_gbs": 6500.0, "bf16_burst": 1650.0, "bf16_sustained": 1400.0, "source": "B200_PROFILING.md fallback"}


def kernel_table(rec, ns, ni, workload):
    """per C-ABI kernel: launches, avg ms, algorithmic work, achieved rate, roofline it is held to."""
    from nerf_fl_b200 import _lib as L
    pk = peaks()
    S_f = ns + ni
    rows = []
    for (name, mode), (n, ms, units) in sorted(rec.items(), key=lambda kv: -kv[1][1]):
        avg_ms = ms / n
        per_launch_units = units / n
        row = {"kernel": name, "mode": mode, "launches": n, "avg_ms": round(avg_ms, 5), "total_ms": round(ms, 4)}
        if name in ("nwb_mlp_fwd", "nwb_mlp_dgrad", "nwb_mlp_wgrad"):
            fine = bool(mode & L.NWB_MLP_TRANSIENT)
            typ = "fine" if fine else ("coarse_sigma" if (mode & L.NWB_MLP_SIGMA_ONLY) else "coarse")
            kind = name.split("_")[-1]
            fl = field_flops_per_point(kind, typ) * per_launch_units
            ach = fl / (avg_ms * 1e-3) / 1e12
            row.update(bound="tensor", achieved=round(ach, 2), peak=pk["bf16_sustained"], unit="TFLOP/s",
                       frac=round(ach / pk["bf16_sustained"], 4), work=f"{typ} {kind}: {field_flops_per_point(kind, typ):.0f} FLOP/point x {per_launch_units:.0f} points")
            if kind == "wgrad":
                # dW = dZ^T . A reads every dZ image and every saved activation image exactly once (DESIGN.md section 3:
                # fine 770 048 + 770 048 B, coarse 638 976 + 638 976 B per 128-point tile) and keeps the products in
                # TMEM: 49 FLOP per byte -- the HBM roofline binds long before the tensor one
                bpp = 12032 if fine else 9984
                gbs = bpp * per_launch_units / (avg_ms * 1e-3) / 1e9
                row.update(bound="hbm", achieved=round(gbs, 1), peak=pk["hbm_gbs"], unit="GB/s", frac=round(gbs / pk["hbm_gbs"], 4),
                           tensor_tflops=round(ach, 2), tensor_frac=round(ach / pk["bf16_sustained"], 4),
                           work=f"{typ} wgrad: {bpp} B/point (dZ + saved activation images, read once) x {per_launch_units:.0f} points")
        elif name in ("nwb_composite_fwd", "nwb_composite_bwd"):
            Cc, S = mode // 1000, mode % 1000
            if name.endswith("fwd"):
                per_ray = S * (4 * Cc + 4 + 4) + (20 if Cc <= 4 else 48)
            else:
                per_ray = S * (4 * Cc + 4 + 4 * Cc) + (32 if Cc <= 4 else 48)
            by = per_ray * per_launch_units
            ach = by / (avg_ms * 1e-3) / 1e9
            row.update(bound="hbm", achieved=round(ach, 1), peak=pk["hbm_gbs"], unit="GB/s", frac=round(ach / pk["hbm_gbs"], 4),
                       work=f"{per_ray} B/ray x {per_launch_units:.0f} rays")
        elif name == "nwb_sample_pdf_merge":
            per_ray = 4 * (ns + ns + S_f)
            by = per_ray * per_launch_units
            ach = by / (avg_ms * 1e-3) / 1e9
            row.update(bound="hbm", achieved=round(ach, 1), peak=pk["hbm_gbs"], unit="GB/s", frac=round(ach / pk["hbm_gbs"], 4),
                       work=f"{per_ray} B/ray x {per_launch_units:.0f} rays")
        elif name == "nwb_sample_coarse":
            per_ray = 32 + 4 * ns
            by = per_ray * per_launch_units
            ach = by / (avg_ms * 1e-3) / 1e9
            row.update(bound="hbm", achieved=round(ach, 1), peak=pk["hbm_gbs"], unit="GB/s", frac=round(ach / pk["hbm_gbs"], 4),
                       work=f"{per_ray} B/ray x {per_launch_units:.0f} rays")
        rows.append(row)
    return rows


def run_product(args):
    import torch
    import torch.distributed as dist
    from nerf_fl_b200 import kernels as K
    from nerf_fl_b200.objective import NerfWLoss
    from nerf_fl_b200.parallel import FlatTrainState, gather_rays, shard_range
    from nerf_fl_b200.rendering import check_errors, render_rays

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback (use --impl reference for the CPU arm)")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    ns = ni = args.samples
    models, emb = build_product(dev, rank)
    lossfn = NerfWLoss()
    hb = host_batches(N_ROTATE, args.batch, rank)
    pinned = [tuple(t.pin_memory() for t in b) for b in hb]
    resident = [tuple(t.to(dev) for t in b) for b in hb]
    loss_host = torch.zeros(1, dtype=torch.float32).pin_memory()

    if args.workload == "train":
        from nerf_fl_b200.train_step import TrainStep
        state = FlatTrainState(list(models.values()) + [emb["a"], emb["t"]], lr=5e-4)
        # the public training-step API: forward, NerfWLoss, backward, NCCL all-reduce of the flat gradient buffer,
        # Adam, zero_grad -- captured once and replayed as a CUDA graph unless --no-graph
        step = TrainStep(models, emb, state, lossfn, batch=args.batch, N_samples=ns, N_importance=ni, perturb=1.0,
                         noise_std=0.0, white_back=True, graph=not args.no_graph)
        eager_step = TrainStep(models, emb, state, lossfn, batch=args.batch, N_samples=ns, N_importance=ni, perturb=1.0,
                               noise_std=0.0, white_back=True, graph=False)
        rays_per_step = args.batch
        h2d = sum(t.numel() * t.element_size() for t in hb[0])
        d2h = 4
    else:
        from nerf_fl_b200 import synthetic as syn
        H = W = 800
        lo, hi = shard_range(H * W, rank, world)
        img = syn.image_rays(H, W)[lo:hi].contiguous()
        img_ts = torch.zeros(hi - lo, dtype=torch.int64)
        pinned = [(img.pin_memory(), img_ts.pin_memory(), None)]
        resident = [(img.to(dev), img_ts.to(dev), None)]
        out_host = torch.zeros(H * W, 4).pin_memory()
        chunk = 32768

        def step(rays, ts, tgt):
            parts = []
            with torch.no_grad():
                for i in range(0, rays.shape[0], chunk):
                    r = render_rays(models, emb, rays[i:i + chunk], ts[i:i + chunk], ns, False, 0, 0, ni, chunk, True,
                                    test_time=True)
                    parts.append(torch.cat([r["rgb_fine"], r["depth_fine"][:, None]], 1))
            return gather_rays(torch.cat(parts, 0), H * W)
        rays_per_step = (hi - lo)
        h2d = img.numel() * 4 + img_ts.numel() * 8
        d2h = H * W * 16
        eager_step = step

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(n_steps, from_host):
        sync_all()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for i in range(n_steps):
            if from_host:
                b = tuple(t.to(dev, non_blocking=True) if t is not None else None for t in pinned[i % len(pinned)])
            else:
                b = resident[i % len(resident)]
            res = step(*b)
            if from_host:
                if args.workload == "train":
                    loss_host.copy_(res.detach().reshape(1), non_blocking=True)
                else:
                    out_host.copy_(res, non_blocking=True)
                torch.cuda.current_stream().synchronize()      # the caller reads the result every step
        e1.record()
        sync_all()
        ms = e0.elapsed_time(e1)
        wall = (time.perf_counter() - t0) * 1e3
        t = torch.tensor([ms, wall], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0]), float(t[1])

    # ---- warm-up, timed region (resident inputs), clocks sampled during it -------------------------
    timed(max(args.warmup, 3), False)
    check_errors()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    l0 = K.launch_count
    ms, wall = timed(args.steps, False)
    launches = K.launch_count - l0
    graphed = args.workload == "train" and getattr(step, "graph", None) is not None
    if graphed:                                  # replays run no Python: count what the captured iteration launched
        launches = step.launches_per_step * args.steps
    clocks = sampler.stop() if rank == 0 else None
    total_rays = rays_per_step * args.steps
    if world > 1:
        tr = torch.tensor([total_rays], device=dev, dtype=torch.float64)
        dist.all_reduce(tr)
        total_rays = float(tr[0])
    value = total_rays / (ms * 1e-3)

    # ---- end to end through the public API from pinned host buffers ---------------------------------
    timed(2, True)
    ms_e2e, _ = timed(args.steps, True)
    e2e_value = total_rays / (ms_e2e * 1e-3)

    # ---- inference only: the same image from a 48-byte pose (rays made on the device) to a uint8 image
    # (clip + pack on the device, 8f-4): what an eval loop built on nerf_fl_b200.inference.render_image moves
    e2e_pose = None
    if args.workload == "infer":
        from nerf_fl_b200.inference import render_image
        Kc, c2w = syn.view_camera(H, W)
        u8_host = torch.zeros(H, W, 3, dtype=torch.uint8).pin_memory()

        def pose_steps(k):
            sync_all()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(k):
                out = render_image(models, emb, H, W, Kc, c2w, 0, ns, ni, chunk=chunk, white_back=True, device=dev)
                u8_host.copy_(out["img_u8"], non_blocking=True)
                torch.cuda.current_stream().synchronize()
            e1.record()
            sync_all()
            t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t[0])
        pose_steps(2)
        ms_pose = pose_steps(args.steps)
        e2e_pose = {"value": H * W * args.steps / (ms_pose * 1e-3), "unit": "rays/s", "h2d_bytes_per_step": 48 + 36,
                    "d2h_bytes_per_step": H * W * 3, "ms_per_step": ms_pose / args.steps,
                    "what": "render_image: pose + intrinsics in (kernel arguments), rays generated on the device, "
                            "rgb_fine clipped and packed to uint8 on the device, (H,W,3) uint8 read back"}

    # ---- per-kernel device times (CUDA events around every C-ABI call, separate pass of the same steps) --
    K.timing_start()
    for i in range(min(args.steps, 10)):
        eager_step(*resident[i % len(resident)])
    rec = K.timing_stop()
    check_errors()
    table = kernel_table(rec, ns, ni, args.workload)
    sum_ms = sum(r["total_ms"] for r in table) / min(args.steps, 10)
    for r in table:
        r["share_of_kernel_time"] = round(r["total_ms"] / min(args.steps, 10) / sum_ms, 4)
    dom = next(r for r in table if "bound" in r)             # table is sorted by total time
    roofline = {"kernel": f"{dom['kernel']} (mode {dom['mode']})", "bound": dom["bound"], "achieved": dom["achieved"],
                "peak": dom["peak"], "unit": dom["unit"], "frac": dom["frac"], "traffic": ncu_traffic(dom, args),
                "traffic_unit": "B per launch (dram read+write, ncu --set full, profiles/ncu_traffic.json)",
                "work_per_launch": dom["work"], "avg_launch_ms": dom["avg_ms"], "peak_source": peaks()["source"] + (" (bf16_tflops_sustained)" if dom["bound"] == "tensor" else " (hbm_gbs, copy read+write)"),
                "timing": "CUDA events around each C-ABI call on the launching stream, separate eager pass of the same steps"}
    fl = (train_flops_per_ray if args.workload == "train" else infer_flops_per_ray)(ns, ni)
    step_roofline = {"flops_per_ray": fl, "achieved_tflops_per_gpu": round(value / world * fl / 1e12, 2),
                     "frac_of_bf16_sustained": round(value / world * fl / 1e12 / peaks()["bf16_sustained"], 4),
                     "sum_kernel_ms_per_step": round(sum_ms, 4), "ms_per_step": round(ms / args.steps, 4)}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    line = {"metric": METRIC, "value": value, "unit": "rays/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak" if args.workload == "train" else "strong",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": workload_config(args, rays_per_step),
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "rays/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": launches, "roofline": roofline, "step_roofline": step_roofline, "kernels": table,
            "wall_ms_per_step": wall / args.steps}
    if e2e_pose is not None:
        line["e2e_pose_in_u8_out"] = e2e_pose
    if world == 1 and not args.no_cpu_baseline:
        v, sample, cores, cms, n = cpu_leg(args.workload, ns, ni, 3, 1, 25.0, min(args.batch, 1024))
        line["cpu_baseline"] = {"value": v, "unit": "rays/s", "cores": cores, "kind": "port", "sample": sample}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


_JSON_FD = None


def emit(line: dict):
    """The ONE JSON line, written to the process's original stdout."""
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    global _JSON_FD, N_VOCAB, BETA_MIN
    args = parse_args()
    N_VOCAB, BETA_MIN = args.vocab, args.beta_min
    # native libraries print to fd 1 behind Python's back (NCCL's "NCCL version ..." banner): keep the real stdout
    # for the JSON line only and send everything else to stderr
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_product(args)


if __name__ == "__main__":
    main()
