#!/usr/bin/env python
"""bench.py -- NeRF-W rays/s on B200 (BASELINE.json metric), one JSON line on stdout.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload train|infer]

Default workload = BASELINE.json configs[1]: a NeRF-W training step (render_rays forward, NerfWLoss,
backward, gradient all-reduce when N>1, Adam on the flat buffer, bf16 operand-cache refresh) on
Blender-shaped synthetic rays, batch 1024 per rank, encode_a + encode_t, beta_min 0.1, 64+64 samples.
`value` is whole-job rays/s with the ray batches resident in HBM; `e2e` is the same step driven from
pinned HOST buffers (H2D of rays/ts/targets and D2H of the loss inside the timed region).

The same JSON line also carries: `infer` -- configs[3], the 800x800 test_time render at chunk
32 768 on the same N GPUs (rgb+depth gathered; select-keys and the reference-faithful all-keys-to-host mode of
eval.py:105-106); `parity_check` -- the loss of one eager product step on batch 0 with pre-drawn RNG tensors against
the CPU oracle on the same inputs; `torch_gpu_baseline` -- the UNMODIFIED reference files (baseline/_ref) moved to
`cuda`, stock ATen, the same training step (section 8d's same-box GPU comparison); `cpu_baseline` (+ `_cfg1`).

`--impl reference` times the reference's OWN files (baseline/_ref/models/{rendering,nerf}.py + losses.py, copied
unmodified by baseline/install_ref.py; `cpu_baseline.kind` = "reference") for the same step on the host cores; if
they are absent it falls back to oracle/nerfw_oracle.py (the pinned torch-CPU restatement, kind "port").  That
fallback, `parity_check` and the port leg are the only places bench.py touches oracle/.
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
    ap.add_argument("--no-infer", action="store_true", help="train workload: skip the configs[3] 800x800 render block")
    ap.add_argument("--no-parity", action="store_true", help="train workload: skip the loss check against the CPU oracle")
    ap.add_argument("--dgrad-gen", type=int, default=0, choices=[0, 1, 2],
                    help="dgrad kernel: 0 = the library's per-launch choice, 1 = first-generation, 2 = two-slot pair kernel (A/B runs)")
    ap.add_argument("--allreduce", default="p2p", choices=["p2p", "p2p-fused-adam", "nccl"],
                    help="N > 1: gradient exchange = our one-kernel all-reduce over NVLink peer memory inside the CUDA graph "
                         "(default) or torch.distributed's NCCL all_reduce behind the graph")
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
def load_reference():
    """(render_rays, NeRF, PosEmbedding, NerfWLoss) of the unmodified reference from baseline/_ref, or None."""
    ref = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.exists(os.path.join(ref, "models", "rendering.py")):
        return None
    if ref not in sys.path:
        sys.path.insert(0, ref)
    try:
        from models.rendering import render_rays as ref_render       # noqa: E402  (baseline/_ref/models)
        from models.nerf import NeRF as RefNeRF, PosEmbedding as RefPos
        from losses import NerfWLoss as RefLoss
    except Exception as e:                                            # noqa: BLE001
        sys.stderr.write(f"bench.py: baseline/_ref does not import ({e}); falling back to the oracle port\n")
        return None
    return ref_render, RefNeRF, RefPos, RefLoss


def reference_setup(ref, device, nerfw=True):
    """The reference's own modules holding the SAME seeded weights as the product arm (state_dict names match)."""
    import torch
    from nerf_fl_b200 import synthetic as syn
    _, RefNeRF, RefPos, RefLoss = ref
    coarse = RefNeRF("coarse")
    coarse.load_state_dict(syn.make_field_params(11, "coarse"))
    if nerfw:
        fine = RefNeRF("fine", encode_appearance=True, encode_transient=True, beta_min=BETA_MIN)
        fine.load_state_dict(syn.make_field_params(12, "fine", True, True))
    else:
        fine = RefNeRF("fine")
        fine.load_state_dict(syn.make_field_params(12, "fine"))
    models = {"coarse": coarse.to(device), "fine": fine.to(device)}
    emb = {"xyz": RefPos(9, 10), "dir": RefPos(3, 4)}
    if nerfw:
        emb["a"] = torch.nn.Embedding.from_pretrained(syn.make_embedding(13, N_VOCAB, 48), freeze=False).to(device)
        emb["t"] = torch.nn.Embedding.from_pretrained(syn.make_embedding(14, N_VOCAB, 16), freeze=False).to(device)
    params = [p for m in models.values() for p in m.parameters()]
    params += [e.weight for k, e in emb.items() if k in ("a", "t")]
    opt = torch.optim.Adam(params, lr=5e-4, eps=1e-8)                 # utils/__init__.py:31
    return models, emb, opt, RefLoss()


def reference_train_step(ref, st, rays, ts, tgt, ns, ni):
    """train.py:99-111,158-174 with the reference's own render_rays / NerfWLoss, then Adam."""
    models, emb, opt, lossfn = st
    opt.zero_grad(set_to_none=True)
    out = ref[0](models, emb, rays, ts, ns, False, 1.0, 0.0, ni, 32 * 1024, True)
    loss = sum(lossfn(out, tgt).values())
    loss.backward()
    opt.step()
    return loss.detach()


def reference_infer_step(ref, st, rays, ts, ns, ni):
    import torch
    models, emb, _, _ = st
    with torch.no_grad():
        out = ref[0](models, emb, rays, ts, ns, False, 0, 0, ni, 32 * 1024, True, test_time=True)
    return out["rgb_fine"]


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
    whole leg fits `budget_s`.  -> (rays/s, sample description, cores, ms/step, rays per step, kind).
    workload: "train" (configs[1]-shaped NeRF-W step), "infer" (NeRF-W test_time render) or "cfg1" (configs[0]: plain
    NeRF test_time forward, no a/t, under no_grad -- the reference's own CPU-runnable case)."""
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    ref = load_reference()
    kind = "reference" if ref is not None else "port"
    rays, ts, tgt = host_batches(1, batch, 0)[0]
    n = batch
    if ref is not None:
        state = reference_setup(ref, "cpu", nerfw=(workload != "cfg1"))
    else:
        state = cpu_train_setup(batch)

    def one(n):
        if ref is not None:
            if workload == "train":
                reference_train_step(ref, state, rays[:n], ts[:n], tgt[:n], ns, ni)
            else:
                reference_infer_step(ref, state, rays[:n], ts[:n] if workload == "infer" else None, ns, ni)
        elif workload == "train":
            cpu_train_step(state, rays[:n], ts[:n], tgt[:n], ns, ni)
        elif workload == "infer":
            cpu_infer_step(state, rays[:n], ts[:n], ns, ni)
        else:
            cpu_cfg1_step(rays[:n], ns, ni)

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
    what = {"train": "NeRF-W training step (render_rays fwd, NerfWLoss, bwd, Adam)", "infer": "NeRF-W test_time render_rays",
            "cfg1": "plain-NeRF test_time render_rays (configs[0]: no encode_a/t, no_grad)"}[workload]
    src = "baseline/_ref (the reference's own models/rendering.py, models/nerf.py, losses.py)" if ref is not None \
        else "oracle/nerfw_oracle.py port"
    sample = (f"{steps} x {n}-ray {what}, {ns}+{ni} samples, after {warmup} warm-up, {src}, torch {torch.__version__} CPU, "
              f"{cores} threads")
    return n * steps / dt, sample, cores, dt / steps * 1e3, n, kind


def cpu_cfg1_step(rays, ns, ni):
    """configs[0] through the oracle port (only when baseline/_ref is absent)."""
    import torch
    from nerf_fl_b200 import synthetic as syn
    from oracle import nerfw_oracle as O
    pc, pf = syn.make_field_params(11, "coarse"), syn.make_field_params(12, "fine")
    with torch.no_grad():
        out = O.render_rays(pc, pf, None, None, rays, None, n_samples=ns, n_importance=ni, perturb=0.0, noise_std=0.0,
                            white_back=True, test_time=True)
    return float(out["rgb_fine"].sum())


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    ns = ni = args.samples
    v, sample, cores, ms, n, kind = cpu_leg(args.workload, ns, ni, args.steps, args.warmup, 150.0, args.batch)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "rays/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, n_rays_step=n),
            "cpu_baseline": {"value": v, "unit": "rays/s", "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": v, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


def ncu_traffic(kernel: str, mode: int, batch: int, samples: int, workload: str):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of a kernel, from the committed `ncu --set full` capture
    of this same command (profiles/ncu_traffic.json, written by tools/ncu_table.py); None when no capture exists for
    this kernel / workload size."""
    try:
        tab = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
    except (OSError, ValueError):
        return None
    ent = tab.get(f"{kernel}|mode{mode}|batch{batch}|samples{samples}|{workload}")
    return None if ent is None else ent["bytes"]          # bytes per launch


def workload_config(args, n_rays_step=None):
    ns = args.samples
    if args.workload == "train":
        name = (f"configs[1]: NeRF-W training step fwd+bwd (+all-reduce, Adam), Blender-shaped synthetic rays 400x400, "
                f"batch {args.batch}/rank, encode_a+encode_t, N_vocab {N_VOCAB}, beta_min {BETA_MIN}, {ns}+{ns} samples, "
                f"perturb 1, noise_std 0, white_back")
    else:
        name = (f"configs[3]: NeRF-W test_time render_rays (eval.py path), 800x800 synthetic image, chunk 32768, {ns}+{ns} samples, "
                f"ray-sharded across ranks, rgb+depth gathered")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    cfg = {"workload": name, "rays_per_step_per_rank": n_rays_step if n_rays_step is not None else args.batch,
           "n_samples": ns, "n_importance": ns,
           "gradient_exchange": ("none (1 GPU)" if world == 1 or args.workload != "train" else
                                 ("nwb_p2p_allreduce: one kernel per rank over NVLink peer memory (CUDA IPC), inside the graph"
                                  if args.allreduce == "p2p" else "NCCL all_reduce of the flat gradient buffer behind the graph")),
           "launch": "eager" if (args.no_graph or args.workload != "train") else "CUDA graph replay of the whole iteration",
           "l2": f"{N_ROTATE} distinct resident ray batches rotate; per-step working set (saved activation + dZ images, "
                 f">1 GB at batch 1024) exceeds the 126 MB L2"}
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


def field_image_bytes_per_point(kind: str, fine: bool):
    """HBM bytes per sample point the field kernels move BY DESIGN (saved activation images, dZ
    images, sign masks) -- implementation traffic, not algorithmic bytes; reported as the `hbm_view` of a tensor-bound
    kernel so that the distance between its tensor fraction and 1 is explained."""
    from nerf_fl_b200 import kernels as K
    return K.field_traffic_bytes_per_point(kind, fine)


def kernel_table(rec, ns, ni, workload):
    """per C-ABI kernel: launches, avg ms, algorithmic work, achieved rate, roofline it is held to (the
    MLP kernels to the TENSOR roofline, compositing / sampling to HBM)."""
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
                       frac=round(ach / pk["bf16_sustained"], 4), frac_burst=round(ach / pk["bf16_burst"], 4),
                       work=f"{typ} {kind}: {field_flops_per_point(kind, typ):.0f} FLOP/point x {per_launch_units:.0f} points")
            bpp = field_image_bytes_per_point(kind, fine) if (kind != "fwd" or (mode & L.NWB_MLP_SAVE)) else 0
            if bpp:
                gbs = bpp * per_launch_units / (avg_ms * 1e-3) / 1e9
                row["hbm_view"] = {"design_bytes_per_point": bpp, "achieved": round(gbs, 1), "unit": "GB/s",
                                   "frac": round(gbs / pk["hbm_gbs"], 4)}
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


def roofline_of(table, n_passes, value_per_gpu, flops_per_ray, ms_per_step, batch, samples, workload):
    """The `roofline` object: the dominant kernel BY TIME held to the roofline SURVEY 8d assigns it (tensor for the
    MLP kernels), with the HBM view of the same launch and the step-level figure beside it."""
    pk = peaks()
    sum_ms = sum(r["total_ms"] for r in table) / n_passes
    for r in table:
        r["share_of_kernel_time"] = round(r["total_ms"] / n_passes / sum_ms, 4)
    dom = next(r for r in table if "bound" in r)             # table is sorted by total time
    roof = {"kernel": f"{dom['kernel']} (mode {dom['mode']})", "bound": dom["bound"], "achieved": dom["achieved"],
            "peak": dom["peak"], "unit": dom["unit"], "frac": dom["frac"],
            "work_per_launch": dom["work"], "avg_launch_ms": dom["avg_ms"],
            "peak_source": pk["source"] + (" (bf16_tflops_sustained; frac_burst vs bf16_tflops)" if dom["bound"] == "tensor"
                                           else " (hbm_gbs, copy read+write)"),
            "timing": "CUDA events around each C-ABI call on the launching stream, separate eager pass of the same steps"}
    if "frac_burst" in dom:
        roof["frac_burst"] = dom["frac_burst"]
    traffic = ncu_traffic(dom["kernel"], dom["mode"], batch, samples, workload)
    roof["traffic"] = traffic
    roof["traffic_unit"] = "B per launch (dram read+write, ncu --set full, profiles/ncu_traffic.json)"
    if "hbm_view" in dom:
        hv = dict(dom["hbm_view"])
        hv["traffic"] = traffic
        hv["traffic_over_design_bytes"] = (round(traffic / (hv["design_bytes_per_point"] * float(dom["work"].split(" x ")[1].split()[0])), 3)
                                           if traffic else None)
        roof["hbm_view"] = hv
    ach = value_per_gpu * flops_per_ray / 1e12
    roof["step"] = {"flops_per_ray": flops_per_ray, "achieved_tflops_per_gpu": round(ach, 2),
                    "frac_of_bf16_sustained": round(ach / pk["bf16_sustained"], 4),
                    "frac_of_bf16_burst": round(ach / pk["bf16_burst"], 4),
                    "sum_kernel_ms_per_step": round(sum_ms, 4), "ms_per_step": round(ms_per_step, 4)}
    return roof


class Ctx:
    pass


def make_ctx(args):
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback (use --impl reference for the CPU arm)")
    c = Ctx()
    c.world = int(os.environ.get("WORLD_SIZE", "1"))
    c.rank = int(os.environ.get("RANK", "0"))
    c.local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(c.local)
    c.dev = torch.device("cuda", c.local)
    if c.world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=c.dev)
    c.ns = c.ni = args.samples
    c.models, c.emb = build_product(c.dev, c.rank)

    def sync_all():
        if c.world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(*vals):
        t = torch.tensor(list(vals), device=c.dev, dtype=torch.float64)
        if c.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(x) for x in t]

    c.sync_all, c.max_over_ranks = sync_all, max_over_ranks
    return c


def measure_infer(c, steps, warmup, with_all_keys=True, with_pose=True):
    """configs[3] (eval.py:76-106,186-205): one 800x800 view rendered with test_time=True in 32 768-ray chunks, rays
    block-sharded over the ranks, rgb+depth gathered.  -> dict with `value` (rays resident), `e2e` (rays from pinned host
    memory, the gathered image read back), `all_keys` (every result key of every chunk copied to the host as the
    reference's batched_inference does), `e2e_pose_in_u8_out` and the dominant kernel's roofline."""
    import torch
    from nerf_fl_b200 import kernels as K
    from nerf_fl_b200 import synthetic as syn
    from nerf_fl_b200.parallel import gather_rays, shard_range
    from nerf_fl_b200.rendering import check_errors, render_rays
    H = W = 800
    chunk = 32768
    ns, ni, dev = c.ns, c.ni, c.dev
    lo, hi = shard_range(H * W, c.rank, c.world)
    img = syn.image_rays(H, W)[lo:hi].contiguous()
    img_ts = torch.zeros(hi - lo, dtype=torch.int64)
    pinned = (img.pin_memory(), img_ts.pin_memory())
    resident = (img.to(dev), img_ts.to(dev))
    out_host = torch.zeros(H * W, 4).pin_memory()

    def render(rays, ts, sink=None):
        parts = []
        with torch.no_grad():
            for i in range(0, rays.shape[0], chunk):
                r = render_rays(c.models, c.emb, rays[i:i + chunk], ts[i:i + chunk], ns, False, 0, 0, ni, chunk, True,
                                test_time=True)
                if sink is not None:
                    sink(i, r)
                parts.append(torch.cat([r["rgb_fine"], r["depth_fine"][:, None]], 1))
        return gather_rays(torch.cat(parts, 0), H * W)

    def timed(n_steps, from_host, sink=None):
        c.sync_all()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n_steps):
            b = tuple(t.to(dev, non_blocking=True) for t in pinned) if from_host else resident
            res = render(*b, sink=sink)
            if from_host:
                out_host.copy_(res, non_blocking=True)
                torch.cuda.current_stream().synchronize()
        e1.record()
        c.sync_all()
        return c.max_over_ranks(e0.elapsed_time(e1))[0]

    timed(max(warmup, 1), False)
    check_errors()
    sampler = ClockSampler(c.local)
    if c.rank == 0:
        sampler.start()
        time.sleep(0.25)
    l0 = K.launch_count
    ms = timed(steps, False)
    launches = K.launch_count - l0
    clocks = sampler.stop() if c.rank == 0 else None
    n_rays = H * W
    out = {"value": n_rays * steps / (ms * 1e-3), "unit": "rays/s", "ms_per_image": ms / steps, "steps": steps,
           "rays_per_image": n_rays, "rays_per_rank": hi - lo, "chunk": chunk, "gpu_launches": launches, "clocks": clocks,
           "keys": "rgb_fine + depth_fine kept on the device and gathered (eval.py:105-106 restricted to the keys eval.py uses)"}
    timed(1, True)
    ms_e2e = timed(steps, True)
    out["e2e"] = {"value": n_rays * steps / (ms_e2e * 1e-3), "unit": "rays/s", "ms_per_image": ms_e2e / steps,
                  "h2d_bytes_per_step": img.numel() * 4 + img_ts.numel() * 8, "d2h_bytes_per_step": H * W * 16}
    if with_all_keys:
        # the reference-faithful mode: EVERY key of every chunk goes to the host (eval.py:105-106 `v.cpu()`)
        hostbuf = {}
        nbytes = [0]

        def sink(i, r):
            for k, v in r.items():
                key = (i, k)
                if key not in hostbuf:
                    hostbuf[key] = torch.empty(v.shape, dtype=v.dtype).pin_memory()
                hostbuf[key].copy_(v, non_blocking=True)
                nbytes[0] += v.numel() * v.element_size()
        timed(1, True, sink)
        nbytes[0] = 0
        ms_all = timed(steps, True, sink)
        out["all_keys"] = {"value": n_rays * steps / (ms_all * 1e-3), "unit": "rays/s", "ms_per_image": ms_all / steps,
                           "d2h_bytes_per_step": nbytes[0] // steps + H * W * 16,
                           "what": "every result key of every chunk copied to pinned host memory, as the reference's "
                                   "batched_inference does (eval.py:105-106), plus the gathered rgb+depth"}
        del hostbuf
    if with_pose:
        from nerf_fl_b200.inference import render_image
        Kc, c2w = syn.view_camera(H, W)
        u8_host = torch.zeros(H, W, 3, dtype=torch.uint8).pin_memory()

        def pose_steps(k):
            c.sync_all()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(k):
                o = render_image(c.models, c.emb, H, W, Kc, c2w, 0, ns, ni, chunk=chunk, white_back=True, device=dev)
                u8_host.copy_(o["img_u8"], non_blocking=True)
                torch.cuda.current_stream().synchronize()
            e1.record()
            c.sync_all()
            return c.max_over_ranks(e0.elapsed_time(e1))[0]
        pose_steps(1)
        ms_pose = pose_steps(steps)
        out["e2e_pose_in_u8_out"] = {"value": n_rays * steps / (ms_pose * 1e-3), "unit": "rays/s", "h2d_bytes_per_step": 48 + 36,
                                     "d2h_bytes_per_step": H * W * 3, "ms_per_image": ms_pose / steps,
                                     "what": "render_image: pose + intrinsics in (kernel arguments), rays generated on the "
                                             "device, rgb_fine clipped and packed to uint8 on the device, (H,W,3) uint8 read back"}
    K.timing_start()
    render(*resident)
    rec = K.timing_stop()
    table = kernel_table(rec, ns, ni, "infer")
    out["roofline"] = roofline_of(table, 1, out["value"] / c.world, infer_flops_per_ray(ns, ni), ms / steps, 0, ns, "infer")
    out["kernels"] = table
    return out


def parity_check(c, args):
    """One EAGER product step on batch 0 with pre-drawn RNG tensors (so product and oracle sample the same depths)
    against the CPU oracle on the same inputs and the same initial weights: the loss the timed configuration computes is
    the reference's.  Runs before any optimizer step; leaves the model untouched."""
    import torch
    from nerf_fl_b200 import synthetic as syn
    from nerf_fl_b200.rendering import check_errors, render_rays
    from oracle import nerfw_oracle as O
    ns, ni, n = c.ns, c.ni, args.batch
    rays, ts, tgt = host_batches(1, n, 0)[0]
    gen = torch.Generator().manual_seed(4242)
    rng = {"perturb_rand": torch.rand(n, ns, generator=gen), "u": torch.rand(n, ni, generator=gen)}
    with torch.no_grad():
        out = render_rays(c.models, c.emb, rays.to(c.dev), ts.to(c.dev), ns, False, 1.0, 0.0, ni, 32768, True,
                          rng={k: v.to(c.dev) for k, v in rng.items()})
        loss_gpu = float(sum(O.nerfw_loss({k: v.float().cpu() for k, v in out.items()}, tgt).values()))
        check_errors()
        t0 = time.perf_counter()
        want = O.render_rays(syn.make_field_params(11, "coarse"), syn.make_field_params(12, "fine", True, True),
                             syn.make_embedding(13, N_VOCAB, 48), syn.make_embedding(14, N_VOCAB, 16), rays, ts,
                             n_samples=ns, n_importance=ni, perturb=1.0, noise_std=0.0, white_back=True, encode_a=True,
                             encode_t=True, beta_min=BETA_MIN, rng=rng)
        loss_cpu = float(sum(O.nerfw_loss(want, tgt).values()))
    worst = {k: float((out[k].float().cpu() - want[k]).abs().max()) for k in ("rgb_fine", "rgb_coarse", "depth_fine", "weights_fine")}
    return {"loss_gpu": loss_gpu, "loss_cpu": loss_cpu, "rel": abs(loss_gpu - loss_cpu) / abs(loss_cpu),
            "max_abs_err": worst, "rays": n, "oracle_s": round(time.perf_counter() - t0, 2),
            "what": "eager render_rays on batch 0 with pre-drawn perturb_rand / u vs oracle/nerfw_oracle.py (fp32 CPU), "
                    "NerfWLoss of both evaluated by the oracle's loss on the host"}


def torch_gpu_baseline(c, args, steps=10, warmup=3):
    """SURVEY 8d 'stock PyTorch on B200': the UNMODIFIED reference files moved to `cuda` -- the reference's render_rays,
    NeRF, NerfWLoss, torch.optim.Adam, ATen / cuBLAS kernels, eager -- on the same training step and batches."""
    import torch
    ref = load_reference()
    if ref is None:
        return {"unavailable": "baseline/_ref not present"}
    ns, ni, n = c.ns, c.ni, args.batch
    batches = [tuple(t.to(c.dev) for t in b) for b in host_batches(2, n, 0)]
    out = {}
    for tf32 in (False, True):
        torch.backends.cuda.matmul.allow_tf32 = tf32
        torch.backends.cudnn.allow_tf32 = tf32
        st = reference_setup(ref, c.dev)
        for i in range(warmup):
            reference_train_step(ref, st, *batches[i % 2], ns, ni)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            loss = reference_train_step(ref, st, *batches[i % 2], ns, ni)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        out["tf32" if tf32 else "fp32"] = {"value": n / (ms * 1e-3), "unit": "rays/s", "ms_per_step": ms, "loss": float(loss)}
        del st
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    torch.cuda.empty_cache()
    out.update(value=out["fp32"]["value"], ms_per_step=out["fp32"]["ms_per_step"], unit="rays/s", steps=steps,
               what=f"baseline/_ref (reference render_rays + NeRF + NerfWLoss, unmodified) on cuda:0, eager ATen, torch "
                    f"{torch.__version__}, batch {n}, {ns}+{ni}, fwd+bwd+Adam; `value` = torch default (fp32 matmul), "
                    f"`tf32` = allow_tf32")
    return out


def run_product(args):
    import torch
    import torch.distributed as dist
    from nerf_fl_b200 import kernels as K
    from nerf_fl_b200.objective import NerfWLoss
    from nerf_fl_b200.parallel import FlatTrainState
    from nerf_fl_b200.rendering import check_errors

    c = make_ctx(args)
    if args.dgrad_gen:
        K.mlp_dgrad_select(args.dgrad_gen)
    world, rank, dev, ns, ni = c.world, c.rank, c.dev, c.ns, c.ni
    warmup = max(args.warmup, 3)

    if args.workload == "infer":
        res = measure_infer(c, args.steps, warmup)
        if rank == 0:
            line = {"metric": METRIC, "value": res["value"], "unit": "rays/s", "n_gpus": world, "steps": args.steps,
                    "warmup": warmup, "ms_per_step": res["ms_per_image"], "higher_is_better": True, "scaling": "strong",
                    "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                    "config": workload_config(args, res["rays_per_rank"]), "clocks": res["clocks"], "e2e": res["e2e"],
                    "gpu_launches": res["gpu_launches"], "roofline": res["roofline"], "kernels": res["kernels"],
                    "all_keys": res.get("all_keys"), "e2e_pose_in_u8_out": res.get("e2e_pose_in_u8_out")}
            if world == 1 and not args.no_cpu_baseline:
                v, sample, cores, cms, n, kind = cpu_leg("infer", ns, ni, 3, 1, 25.0, 1024)
                line["cpu_baseline"] = {"value": v, "unit": "rays/s", "cores": cores, "kind": kind, "sample": sample}
            emit(line)
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- training workload (configs[1]) --------------------------------------------------------------
    lossfn = NerfWLoss()
    par = parity_check(c, args) if (rank == 0 and not args.no_parity) else None     # before any optimizer step
    if world > 1:
        dist.barrier()
    hb = host_batches(N_ROTATE, args.batch, rank)
    pinned = [tuple(t.pin_memory() for t in b) for b in hb]
    resident = [tuple(t.to(dev) for t in b) for b in hb]
    loss_host = torch.zeros(1, dtype=torch.float32).pin_memory()
    from nerf_fl_b200.train_step import TrainStep
    state = FlatTrainState(list(c.models.values()) + [c.emb["a"], c.emb["t"]], lr=5e-4)
    if world > 1 and args.allreduce == "p2p-fused-adam":       # A/B: reduce-scatter kernel + optimizer reading the reduced slices from the peers
        state.p2p_fused_adam = True
        args.allreduce = "p2p"
    if world > 1 and args.allreduce == "p2p" and not state.enable_p2p():
        args.allreduce = "nccl"                        # peer mapping failed on some rank (all ranks agree): reported as the NCCL arm
    # the public training-step API: forward, NerfWLoss, backward, NCCL all-reduce of the flat gradient buffer,
    # Adam, zero_grad -- captured once and replayed as a CUDA graph unless --no-graph
    step = TrainStep(c.models, c.emb, state, lossfn, batch=args.batch, N_samples=ns, N_importance=ni, perturb=1.0,
                     noise_std=0.0, white_back=True, graph=not args.no_graph)
    eager_step = TrainStep(c.models, c.emb, state, lossfn, batch=args.batch, N_samples=ns, N_importance=ni, perturb=1.0,
                           noise_std=0.0, white_back=True, graph=False)
    h2d = sum(t.numel() * t.element_size() for t in hb[0])
    d2h = 4

    def timed(n_steps, from_host):
        c.sync_all()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for i in range(n_steps):
            if from_host:
                b = pinned[i % len(pinned)]             # TrainStep copies pinned host batches straight into its static buffers
            else:
                b = resident[i % len(resident)]
            res = step(*b)
            if from_host:
                loss_host.copy_(res.detach().reshape(1), non_blocking=True)
                torch.cuda.current_stream().synchronize()      # the caller reads the loss every step
        e1.record()
        c.sync_all()
        return c.max_over_ranks(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3)

    # ---- warm-up, timed region (resident inputs), clocks sampled during it -------------------------
    timed(warmup, False)
    check_errors()
    sampler = ClockSampler(c.local)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    l0 = K.launch_count
    ms, wall = timed(args.steps, False)
    launches = K.launch_count - l0
    if getattr(step, "graph", None) is not None:     # replays run no Python: count what the captured iteration launched
        launches = step.launches_per_step * args.steps
    clocks = sampler.stop() if rank == 0 else None
    total_rays = float(args.batch * args.steps * world)
    value = total_rays / (ms * 1e-3)

    # ---- end to end through the public API from pinned host buffers ---------------------------------
    timed(2, True)
    ms_e2e, _ = timed(args.steps, True)
    e2e_value = total_rays / (ms_e2e * 1e-3)
    final_loss = float(loss_host[0])

    # ---- per-kernel device times (CUDA events around every C-ABI call, separate eager pass of the same steps) --
    n_pass = min(args.steps, 10)
    K.timing_start()
    for i in range(n_pass):
        eager_step(*resident[i % len(resident)])
    rec = K.timing_stop()
    check_errors()
    table = kernel_table(rec, ns, ni, "train")
    roofline = roofline_of(table, n_pass, value / world, train_flops_per_ray(ns, ni), ms / args.steps, args.batch,
                           args.samples, "train")

    # ---- configs[3] on the same GPUs, in the same line (the driver's SCALE run then carries it at 1/2/4/8) ----
    infer = None
    if not args.no_infer:
        del step, eager_step
        torch.cuda.empty_cache()
        infer = measure_infer(c, min(args.steps, 5), 2, with_all_keys=True, with_pose=True)
        infer.pop("kernels", None)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    line = {"metric": METRIC, "value": value, "unit": "rays/s", "n_gpus": world, "steps": args.steps, "warmup": warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": workload_config(args, args.batch),
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "rays/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": launches, "roofline": roofline, "kernels": table, "wall_ms_per_step": wall / args.steps,
            "loss_after_timed_steps": final_loss}
    if par is not None:
        line["parity_check"] = par
    if infer is not None:
        line["infer"] = infer
    if world == 1 and not args.no_cpu_baseline:
        line["torch_gpu_baseline"] = torch_gpu_baseline(c, args)
        v, sample, cores, cms, n, kind = cpu_leg("train", ns, ni, 3, 1, 25.0, min(args.batch, 1024))
        line["cpu_baseline"] = {"value": v, "unit": "rays/s", "cores": cores, "kind": kind, "sample": sample}
        v1, sample1, cores1, _, _, kind1 = cpu_leg("cfg1", 64, 64, 3, 1, 10.0, 1024)
        line["cpu_baseline_cfg1"] = {"value": v1, "unit": "rays/s", "cores": cores1, "kind": kind1, "sample": sample1}
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
