"""Data-parallel plumbing for the ray path: rays are independent, so the only
exchange step of a training iteration is the gradient all-reduce that Lightning's DDP performs for
the reference (train.py:235-236), and inference shards each image into contiguous ray blocks with
one final gather (the eval.py:186-194 loop, sharded).

``FlatTrainState`` re-homes every trainable tensor of the NeRF models and embedding tables into ONE
flat fp32 parameter buffer and ONE flat gradient buffer.  The backward kernels accumulate straight
into the gradient views (no per-tensor autograd traffic), so a single in-place NCCL all-reduce on
the flat buffer -- enqueued on the compute stream right behind the wgrad kernel -- is the whole
collective, and ``nwb_adam_step`` (csrc/optim.cu) is the whole optimizer.  One process per GPU,
torch.distributed for the plumbing (NCCL on GPUs; gloo in the CPU tests of the host logic).
"""
from __future__ import annotations

import ctypes as C
import math

import torch
import torch.distributed as dist

__all__ = ["FlatTrainState", "FlatAdam", "P2PGradExchange", "shard_range", "gather_rays", "dist_info"]


def dist_info():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_range(n: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous block partition of n rays: rank r owns [lo, hi); sizes differ by at most one."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def gather_rays(local: torch.Tensor, n_total: int, group=None) -> torch.Tensor:
    """All-gather of per-ray results (n_local, C) from block-sharded ranks -> (n_total, C) on every
    rank (16 B/ray for rgb+depth: 10 MB per 800x800 image).  Shards are padded to the largest block so
    one all_gather_into_tensor moves everything."""
    rank, world = dist_info()
    if world == 1:
        return local
    width = local.shape[1:]
    per = math.ceil(n_total / world)
    pad = torch.zeros((per, *width), dtype=local.dtype, device=local.device)
    pad[:local.shape[0]] = local
    out = torch.empty((world * per, *width), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, pad, group=group)
    parts = []
    for r in range(world):
        lo, hi = shard_range(n_total, r, world)
        parts.append(out[r * per:r * per + (hi - lo)])
    return torch.cat(parts, 0)


class _DevArray:
    """A raw device pointer as something torch.as_tensor can wrap without copying."""

    def __init__(self, ptr: int, numel: int):
        self.__cuda_array_interface__ = {"shape": (int(numel),), "typestr": "<f4", "data": (int(ptr), False), "version": 2}


class P2PGradExchange:
    """The gradient all-reduce as ONE kernel per rank over NVLink / NVSwitch peer memory (csrc/p2p.cu): every rank's flat
    gradient buffer lives in a cudaMalloc'ed region that all other ranks map through CUDA IPC; ``allreduce()`` launches
    ``nwb_p2p_allreduce`` on the current stream -- no NCCL call, no host synchronisation, so it can be captured inside the
    training iteration's CUDA graph (``TrainStep``).  ``torch.distributed`` is used once, to exchange the 64-byte IPC
    handles.  Sums are taken in rank order on every rank: all ranks end with bit-identical gradients."""

    def __init__(self, numel: int, device: torch.device, group=None):
        from . import _lib as L
        rank, world = dist_info()
        if world < 2 or device.type != "cuda":
            raise RuntimeError("P2PGradExchange needs an initialised process group of >= 2 ranks and CUDA buffers")
        lib = L.lib()
        self.rank, self.world, self.numel, self.device = rank, world, int(numel), device
        self._lib = lib
        lib.nwb_p2p_region_bytes.restype = C.c_size_t
        lib.nwb_p2p_region_bytes.argtypes = [C.c_int64, C.c_int]
        lib.nwb_p2p_grad_offset.restype = C.c_size_t
        lib.nwb_p2p_alloc.argtypes = [C.c_size_t, C.POINTER(C.c_void_p)]
        lib.nwb_p2p_export.argtypes = [C.c_void_p, C.c_char_p]
        lib.nwb_p2p_open.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
        lib.nwb_p2p_close.argtypes = [C.c_void_p]
        lib.nwb_p2p_free.argtypes = [C.c_void_p]
        lib.nwb_p2p_allreduce.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int64, C.c_void_p, C.c_void_p]
        lib.nwb_p2p_reduce_scatter.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int64, C.c_void_p, C.c_void_p]
        nbytes = lib.nwb_p2p_region_bytes(self.numel, world)
        self._base, self._opened, err = None, [], None
        with torch.cuda.device(device):
            # every step below is followed by a collective, so a rank that fails (no peer access, IPC not permitted, out
            # of memory) must keep walking the same sequence: errors are collected and agreed on at the end
            handle = C.create_string_buffer(64)
            try:
                base = C.c_void_p()
                L.check(lib.nwb_p2p_alloc(nbytes, C.byref(base)), "nwb_p2p_alloc")
                self._base = base.value
                L.check(lib.nwb_p2p_export(self._base, handle), "nwb_p2p_export")
            except RuntimeError as e:
                err = e
            handles = [None] * world
            dist.all_gather_object(handles, (handle.raw, err is None), group=group)
            self._regions = (C.c_void_p * world)()
            if err is None and all(ok for _, ok in handles):
                try:
                    for q in range(world):
                        if q == rank:
                            self._regions[q] = self._base
                        else:
                            ptr = C.c_void_p()
                            L.check(lib.nwb_p2p_open(handles[q][0], C.byref(ptr)), f"nwb_p2p_open(rank {q})")
                            self._regions[q] = ptr.value
                            self._opened.append(ptr.value)
                except RuntimeError as e:
                    err = e
            elif err is None:
                err = RuntimeError("another rank could not allocate / export its exchange region")
            ok = torch.tensor([0 if err is not None else 1], dtype=torch.int32, device=device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)      # also: every rank has mapped every region
            if int(ok.item()) == 0:
                self.close(collective=False)
                raise RuntimeError(f"P2PGradExchange: peer mapping failed on at least one rank ({err})")
            self.counter = torch.zeros(4, dtype=torch.int32, device=device)
            self.flat_g = torch.as_tensor(_DevArray(self._base + lib.nwb_p2p_grad_offset(), self.numel), device=device)
            torch.cuda.synchronize(device)

    def allreduce(self) -> None:
        from . import _lib as L
        from . import kernels as K
        with K._timed("nwb_p2p_allreduce", self.numel, dev=self.device):
            L.check(self._lib.nwb_p2p_allreduce(self._regions, self.rank, self.world, self.numel, self.counter.data_ptr(),
                                                L.stream_ptr(self.device)), "nwb_p2p_allreduce")

    def reduce_scatter(self) -> None:
        """First half only: slice q of the summed gradient ends in rank q's reduced buffer (consumer:
        ``nwb_adam_pack_step_p2p``); ``flat_g`` keeps the local gradients."""
        from . import _lib as L
        from . import kernels as K
        with K._timed("nwb_p2p_reduce_scatter", self.numel, dev=self.device):
            L.check(self._lib.nwb_p2p_reduce_scatter(self._regions, self.rank, self.world, self.numel, self.counter.data_ptr(),
                                                     L.stream_ptr(self.device)), "nwb_p2p_reduce_scatter")

    def close(self, collective: bool = True) -> None:
        """Unmap the peers' regions and free the own one (after a barrier: nobody may still be reading it)."""
        if getattr(self, "_base", None) is None and not getattr(self, "_opened", None):
            return
        torch.cuda.synchronize(self.device)
        if collective and dist.is_initialized():
            dist.barrier()
        for ptr in self._opened:
            self._lib.nwb_p2p_close(ptr)
        if self._base is not None:
            self._lib.nwb_p2p_free(self._base)
        self._base, self._opened = None, []


class FlatTrainState:
    """Flat parameter / gradient / Adam-moment buffers over ``modules`` (NeRF models, nn.Embedding).

    After construction every parameter's ``.data`` and ``.grad`` are views into the flat buffers
    (state_dict keys and shapes are untouched, so checkpoints keep loading) and the parameters are
    tagged so the render_rays backward accumulates into ``.grad`` in place."""

    def __init__(self, modules, lr=5e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, fused_pack=True,
                 host_update=None):
        # host_update(state, world, zero_grad): the optimizer update for NON-CUDA buffers.  The product has none (the
        # CUDA kernels run or the call raises); the gloo world-size-2 tests of the sharding / all-reduce plumbing
        # inject a torch-op Adam from tests/host_models.py.
        self._host_update = host_update
        self.modules = list(modules)
        params, seen = [], set()
        for m in self.modules:
            for p in m.parameters():
                if p.requires_grad and id(p) not in seen:
                    seen.add(id(p))
                    params.append(p)
        if not params:
            raise ValueError("FlatTrainState: nothing to train")
        dev = params[0].device
        self.params = params
        # each tensor starts on a 16 B boundary (vectorised kernels, wgrad flush)
        offs, off = [], 0
        for p in params:
            offs.append(off)
            off += (p.numel() + 3) // 4 * 4
        self.numel = off
        self.flat_p = torch.zeros(off, dtype=torch.float32, device=dev)
        self.flat_g = torch.zeros(off, dtype=torch.float32, device=dev)
        self.exp_avg = torch.zeros(off, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(off, dtype=torch.float32, device=dev)
        for p, o in zip(params, offs):
            n = p.numel()
            self.flat_p[o:o + n].copy_(p.data.reshape(-1))
            p.data = self.flat_p[o:o + n].view(p.shape)
            p.grad = self.flat_g[o:o + n].view(p.shape)
            p._nwb_direct_grad = True
        self._offs = offs
        self._p2p = None
        self.betas, self.eps, self.weight_decay = betas, eps, weight_decay
        self.step_count = 0
        # step state of the CUDA optimizer kernel lives on the device ({int32 step, lr, lr/bc1, 1/sqrt(bc2)}), so a
        # captured CUDA graph of the whole iteration can be replayed (nwb_adam_step_dev)
        # (+ the ticket word of nwb_adam_pack_step_dev)
        # 12 words: {int32 step, lr, lr/bc1, 1/sqrt(bc2), ticket, int32 step of the products, double beta1^t, double beta2^t}
        self.hyper = torch.zeros(12, dtype=torch.float32, device=dev) if dev.type == "cuda" else None
        # (f2) fields whose packed operand caches the optimizer kernel refreshes itself: NeRF modules with every
        # parameter inside the flat buffer
        self._fused = None
        self._fused_fields = []
        if fused_pack and dev.type == "cuda":
            mine = {id(p) for p in params}
            self._fused_fields = [m for m in self.modules if hasattr(m, "nwb_packed")
                                  and all(id(p) in mine for p in m.parameters())][:4]
        self._lr = float(lr)
        self.lr = lr
        self.n_params = sum(p.numel() for p in params)

    @property
    def lr(self) -> float:
        return self._lr

    @lr.setter
    def lr(self, value: float) -> None:
        """Schedulers (utils/__init__.py:44-61) set this between steps; stream-ordered device write."""
        self._lr = float(value)
        if self.hyper is not None:
            self.hyper[1:2].fill_(self._lr)

    # ---- collective -------------------------------------------------------------------------
    def enable_p2p(self, group=None) -> bool:
        """Move the flat gradient buffer into a peer-mapped exchange region and switch ``allreduce_grads`` from the NCCL
        call to the one-kernel P2P all-reduce (csrc/p2p.cu).  Call once, after ``init_process_group`` and before the
        first step / graph capture.  Returns False (and changes nothing) with a single rank or CPU buffers."""
        if self._p2p is not None:
            return True
        if dist_info()[1] < 2 or not self.flat_g.is_cuda:
            return False
        try:
            ex = P2PGradExchange(self.numel, self.flat_g.device, group)
        except RuntimeError as e:                        # agreed on by all ranks (see P2PGradExchange): NCCL stays the exchange
            import warnings
            warnings.warn(f"nerf_fl_b200: {e}; falling back to torch.distributed all_reduce")
            return False
        ex.flat_g.copy_(self.flat_g)
        self.flat_g = ex.flat_g
        for p, o in zip(self.params, self._offs):
            p.grad = self.flat_g[o:o + p.numel()].view(p.shape)
        self._p2p = ex
        return True

    @property
    def p2p_enabled(self) -> bool:
        return self._p2p is not None

    def allreduce_grads(self, group=None):
        """Sum over ranks, in place on the flat buffer, stream-ordered behind the wgrad kernels; the
        1/world of DDP's mean is folded into the optimizer kernel's grad_scale."""
        rank, world = dist_info()
        if world > 1:
            if self._p2p is not None:
                self._p2p.allreduce()
            else:
                dist.all_reduce(self.flat_g, op=dist.ReduceOp.SUM, group=group)
        return world

    def zero_grad(self):
        self.flat_g.zero_()

    # ---- optimizer ---------------------------------------------------------------------------
    def adam_step(self, world: int = 1, zero_grad: bool = True, from_slices: bool = False):
        """utils/__init__.py:31 Adam on the flat buffers: one CUDA launch.  No CPU path (see ``host_update``).
        ``from_slices``: the gradients are the ranks' reduced slices (after ``P2PGradExchange.reduce_scatter``), read
        from peer memory by the optimizer kernel itself."""
        self.step_count += 1
        b1, b2 = self.betas
        if from_slices:
            from . import _lib as L
            from . import kernels as K
            fz = self._fused or self._fused_setup()
            ex = self._p2p
            with K._timed("nwb_adam_pack_step_p2p", 0, dev=self.flat_p.device):
                L.check(fz["fn_p2p"](self.flat_p.data_ptr(), self.flat_g.data_ptr(), self.exp_avg.data_ptr(),
                                     self.exp_avg_sq.data_ptr(), self.numel, b1, b2, self.eps, self.weight_decay,
                                     self.hyper.data_ptr(), 1.0 / world, int(zero_grad), fz["n"], fz["descs"], fz["params"],
                                     fz["packed"], fz["rest"].data_ptr(), fz["rest"].numel(), ex._regions, ex.world,
                                     L.stream_ptr(self.flat_p.device)), "nwb_adam_pack_step_p2p")
            for m in self.modules:
                if hasattr(m, "_nwb_cache") and not any(m is f for f in self._fused_fields):
                    m._nwb_dirty = True
            return
        if self.flat_p.is_cuda and self._fused_fields:
            from . import _lib as L
            from . import kernels as K
            fz = self._fused or self._fused_setup()
            early, self._early = getattr(self, "_early", ()), ()
            n_f, descs, params, packed = self._field_tables(fz, [i for i in range(fz["n"]) if i not in early])
            with K._timed("nwb_adam_pack_step_dev", 0, dev=self.flat_p.device):
                L.check(fz["fn"](self.flat_p.data_ptr(), self.flat_g.data_ptr(), self.exp_avg.data_ptr(),
                                 self.exp_avg_sq.data_ptr(), self.numel, b1, b2, self.eps, self.weight_decay,
                                 self.hyper.data_ptr(), 1.0 / world, int(zero_grad), n_f, descs, params,
                                 packed, fz["rest"].data_ptr(), fz["rest"].numel(),
                                 L.stream_ptr(self.flat_p.device)), "nwb_adam_pack_step_dev")
            for m in self.modules:                   # fields outside the fused set still repack on their own
                if hasattr(m, "_nwb_cache") and not any(m is f for f in self._fused_fields):
                    m._nwb_dirty = True
            return
        if self.flat_p.is_cuda:
            from . import _lib as L
            from . import kernels as K
            fn = L.bind("nwb_adam_step_dev", [C.c_void_p] * 4 + [C.c_int64] + [C.c_float] * 4 + [C.c_void_p, C.c_float, C.c_int,
                                                                                               C.c_void_p])
            with K._timed("nwb_adam_step_dev", 0, dev=self.flat_p.device):
                L.check(fn(self.flat_p.data_ptr(), self.flat_g.data_ptr(), self.exp_avg.data_ptr(),
                           self.exp_avg_sq.data_ptr(), self.numel, b1, b2, self.eps, self.weight_decay, self.hyper.data_ptr(),
                           1.0 / world, int(zero_grad), L.stream_ptr(self.flat_p.device)), "nwb_adam_step_dev")
        elif self._host_update is not None:
            self._host_update(self, world, zero_grad)
        else:
            raise RuntimeError("nerf_fl_b200.FlatTrainState: parameters are not on a CUDA device and there is no CPU "
                               "optimizer path")
        for m in self.modules:                       # packed bf16 operand caches are stale now
            if hasattr(m, "_nwb_cache"):
                m._nwb_dirty = True

    @property
    def refreshes_caches(self) -> bool:
        """True when the optimizer launch itself rewrites the packed bf16 operand caches (no repack pass)."""
        return bool(self._fused_fields)

    def prepare(self) -> None:
        """One-time setup of the cache-refreshing optimizer launch; call before capturing a CUDA graph."""
        if self._fused_fields and self._fused is None:
            self._fused_setup()

    def _field_tables(self, fz, idx):
        """(n, descs, params, packed) pointer tables of a subset of the fused fields."""
        idx = tuple(idx)
        if len(idx) == fz["n"]:
            return fz["n"], fz["descs"], fz["params"], fz["packed"]
        sub = fz.setdefault("subsets", {})
        if idx not in sub:
            from . import _lib as L
            n = len(idx)
            descs = (C.POINTER(L.FieldDesc) * n)(*[fz["descs"][i] for i in idx])
            params = (C.POINTER(L.FieldParams) * n)(*[fz["params"][i] for i in idx])
            packed = (C.c_void_p * n)(*[fz["packed"][i] for i in idx])
            sub[idx] = (n, descs, params, packed)
        return sub[idx]

    def adam_step_early(self, modules, zero_grad: bool = True) -> bool:
        """This step's Adam update for some of the fused fields NOW, on the current stream, before the rest of the
        iteration has finished (a model whose backward is complete, while another model's still runs): the same scalars as
        the step's main launch, the step state left where it is (NWB_ADAM_NO_ADVANCE).  The next ``adam_step`` then covers
        the remaining fields and everything else, and advances the state.  One GPU only (with a gradient exchange the
        update waits for it); returns False, doing nothing, when it does not apply."""
        if not (self.flat_p.is_cuda and self._fused_fields) or dist_info()[1] > 1:
            return False
        from . import _lib as L
        from . import kernels as K
        fz = self._fused or self._fused_setup()
        idx = [i for i, f in enumerate(self._fused_fields) if any(f is m for m in modules)]
        if not idx or len(idx) == fz["n"] or getattr(self, "_early", ()):
            return False
        n_f, descs, params, packed = self._field_tables(fz, idx)
        b1, b2 = self.betas
        with K._timed("nwb_adam_pack_step_dev", 0, 1, dev=self.flat_p.device):
            L.check(fz["fn"](self.flat_p.data_ptr(), self.flat_g.data_ptr(), self.exp_avg.data_ptr(),
                             self.exp_avg_sq.data_ptr(), self.numel, b1, b2, self.eps, self.weight_decay,
                             self.hyper.data_ptr(), 1.0, int(zero_grad) | 2, n_f, descs, params, packed, None, 0,
                             L.stream_ptr(self.flat_p.device)), "nwb_adam_pack_step_dev")
        self._early = tuple(idx)
        return True

    def _fused_setup(self):
        """One-time: the packed caches' buffers, the C pointer tables and the `rest` offset list (flat offsets that no
        cache element comes from) for nwb_adam_pack_step_dev."""
        from . import _lib as L
        from . import kernels as K
        fields = self._fused_fields
        n = len(fields)
        keep = []
        descs = (C.POINTER(L.FieldDesc) * n)()
        params = (C.POINTER(L.FieldParams) * n)()
        packed = (C.c_void_p * n)()
        for i, m in enumerate(fields):
            buf = m.nwb_packed()                       # packs once; from here on the optimizer keeps it fresh
            d = m.nwb_desc()
            st, ka = K.field_slots({k: v.detach() for k, v in m.nwb_params().items()})
            keep += [buf, d, st, ka]
            descs[i] = C.pointer(d)
            params[i] = C.pointer(st)
            packed[i] = buf.data_ptr()
        dev = self.flat_p.device
        cover = torch.zeros(self.numel, dtype=torch.uint8, device=dev)
        tables = [C.c_int, C.POINTER(C.POINTER(L.FieldDesc)), C.POINTER(C.POINTER(L.FieldParams)), C.POINTER(C.c_void_p)]
        cov = L.bind("nwb_adam_pack_cover", [C.c_void_p, C.c_int64] + tables + [C.c_void_p, C.c_void_p])
        L.check(cov(self.flat_p.data_ptr(), self.numel, n, descs, params, packed, cover.data_ptr(), L.stream_ptr(dev)),
                "nwb_adam_pack_cover")
        rest = torch.nonzero(cover == 0).flatten().to(torch.int32).contiguous()
        fn = L.bind("nwb_adam_pack_step_dev", [C.c_void_p] * 4 + [C.c_int64] + [C.c_float] * 4 +
                    [C.c_void_p, C.c_float, C.c_int] + tables + [C.c_void_p, C.c_int64, C.c_void_p])
        fn_p2p = L.bind("nwb_adam_pack_step_p2p", [C.c_void_p] * 4 + [C.c_int64] + [C.c_float] * 4 +
                        [C.c_void_p, C.c_float, C.c_int] + tables + [C.c_void_p, C.c_int64, C.POINTER(C.c_void_p), C.c_int, C.c_void_p])
        self._fused = {"fn": fn, "fn_p2p": fn_p2p, "n": n, "descs": descs, "params": params, "packed": packed, "rest": rest,
                       "keep": keep, "covered": int(self.numel - rest.numel())}
        return self._fused

    # ---- checkpoint / resume (train.py:231 resume_from_checkpoint) -------------------------------
    def optimizer_state(self) -> dict:
        """The optimizer state in torch.optim.Adam's own layout ({index: {step, exp_avg, exp_avg_sq}}, per parameter
        in ``self.params`` order), so a Lightning checkpoint of FlatAdam round-trips with one of torch.optim.Adam."""
        out, off = {}, 0
        for i, p in enumerate(self.params):
            n = p.numel()
            out[i] = {"step": torch.tensor(float(self.step_count)),
                      "exp_avg": self.exp_avg[off:off + n].view(p.shape).clone(),
                      "exp_avg_sq": self.exp_avg_sq[off:off + n].view(p.shape).clone()}
            off += (n + 3) // 4 * 4
        return out

    def load_optimizer_state(self, state: dict) -> None:
        off, step = 0, None
        for i, p in enumerate(self.params):
            n = p.numel()
            st = state.get(i, state.get(str(i)))
            if st is not None:
                self.exp_avg[off:off + n].copy_(st["exp_avg"].reshape(-1))
                self.exp_avg_sq[off:off + n].copy_(st["exp_avg_sq"].reshape(-1))
                step = int(st["step"]) if step is None else step
            off += (n + 3) // 4 * 4
        self.step_count = int(step or 0)
        if self.hyper is not None:                   # device-resident step word of nwb_adam_*_step_dev ({int32 step, ...})
            self.hyper[0:1].copy_(torch.tensor([self.step_count], dtype=torch.int32).view(torch.float32))

    def snapshot(self) -> dict:
        """Everything an optimizer step changes (for TrainStep's warm-up, which must not move the model)."""
        snap = {"flat_p": self.flat_p.clone(), "exp_avg": self.exp_avg.clone(), "exp_avg_sq": self.exp_avg_sq.clone(),
                "step_count": self.step_count}
        if self.hyper is not None:
            snap["hyper"] = self.hyper.clone()
        return snap

    def restore(self, snap: dict) -> None:
        self.flat_p.copy_(snap["flat_p"])
        self.exp_avg.copy_(snap["exp_avg"])
        self.exp_avg_sq.copy_(snap["exp_avg_sq"])
        self.step_count = snap["step_count"]
        if self.hyper is not None:
            self.hyper.copy_(snap["hyper"])
        self.flat_g.zero_()
        for m in self.modules:                       # parameters moved back: operand caches must follow
            if hasattr(m, "_nwb_cache"):
                m._nwb_dirty = True

    def pull_lr(self) -> None:
        """Take the learning rate from the FlatAdam facade's param_groups when a torch scheduler moved it."""
        src = getattr(self, "_lr_source", None)
        if src is not None:
            lr = float(src.param_groups[0]["lr"])
            if lr != self._lr:
                self.lr = lr

    def step(self, group=None):
        """all-reduce + Adam: everything that follows backward in one training iteration.  ``p2p_fused_adam = True`` (off by
        default: measured slower) skips the all-gather half of the P2P all-reduce: the reduce-scatter kernel leaves slice q
        of the summed gradient with rank q, and the optimizer kernel reads the slices over NVLink itself."""
        self.pull_lr()
        # [B200, 2 GPUs, same box: profiles/r02z4_exchange_ab.log] full P2P all-reduce kernel + plain optimizer launch 1.147 ms per
        # step; reduce-scatter + optimizer reading the slices from the peers 1.166 ms (its scattered system-scope loads over
        # NVLink cost more than the all-gather half they replace); NCCL behind the graph 1.153 ms; one GPU 1.115 ms.
        if self._p2p is not None and self._fused_fields and self.flat_p.is_cuda and getattr(self, "p2p_fused_adam", False):
            self._p2p.reduce_scatter()
            self.adam_step(self._p2p.world, from_slices=True)
        else:
            self.adam_step(self.allreduce_grads(group))


class FlatAdam(torch.optim.Optimizer):
    """``torch.optim.Optimizer`` facade over a FlatTrainState, so the reference's optimizer / scheduler plumbing
    (utils/__init__.py:24-61: ``get_optimizer`` -> Adam, ``get_scheduler`` -> MultiStepLR | CosineAnnealingLR | LambdaLR,
    optionally wrapped by GradualWarmupScheduler; Lightning's ``configure_optimizers`` train.py:139-142) keeps working:
    schedulers write ``param_groups[0]['lr']``, ``step()`` pushes it to the device-resident step state and runs the
    all-reduce + the one Adam launch.  ``zero_grad`` is a no-op: the kernel clears the flat gradient buffer, and the
    parameters' ``.grad`` views must stay attached to it."""

    def __init__(self, state: FlatTrainState):
        self.flat = state
        super().__init__(state.params, dict(lr=state.lr, betas=state.betas, eps=state.eps,
                                            weight_decay=state.weight_decay))
        state._lr_source = self

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        self.flat.step()
        return loss

    def zero_grad(self, set_to_none: bool = False):
        return None

    def state_dict(self):
        """torch.optim.Adam-compatible: the moments and step live in the flat buffers, not in ``self.state``."""
        sd = super().state_dict()
        sd["state"] = self.flat.optimizer_state()
        return sd

    def load_state_dict(self, state_dict):
        groups = state_dict.get("param_groups")
        if groups:
            for g, src in zip(self.param_groups, groups):
                g.update({k: v for k, v in src.items() if k != "params"})
            self.flat.lr = float(self.param_groups[0]["lr"])
        self.flat.load_optimizer_state(state_dict.get("state", {}))
