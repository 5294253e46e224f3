"""One training iteration of the reference's ``NeRFSystem.training_step`` (train.py:158-174) plus what
Lightning runs around it -- ``loss.backward()``, the DDP gradient all-reduce and ``Adam.step`` -- as ONE
replayable CUDA graph.

At the reference's batch size (1024 rays) an iteration is ~60 kernels of 2-400 us each; launched one by
one from Python the host cannot keep the GPU fed.  ``TrainStep`` runs the iteration eagerly a few times,
captures it (``torch.cuda.graph``) on static input buffers and replays it afterwards: no Python, no
allocator, no launch gaps in the steady state.  Everything inside is stream-ordered device work, so
the capture needs nothing special: random draws come from torch's graph-safe Philox generator, the
optimizer's step counter lives on the device (``nwb_adam_step_dev``), the bf16 operand caches are
refreshed by the captured pack kernels at the start of every replay.
"""
from __future__ import annotations

import torch

from . import kernels as K
from .parallel import FlatTrainState, dist_info
from .rendering import render_rays

__all__ = ["TrainStep"]


class TrainStep:
    """``step(rays, ts, rgbs) -> loss``: render_rays forward, loss, backward, all-reduce, Adam.

    ``models`` / ``embeddings`` as for ``render_rays``; ``state`` a ``FlatTrainState`` over them;
    ``loss_fn(results, rgbs) -> dict`` (``NerfWLoss``).  ``batch`` fixes the ray count of the captured
    graph; other batch sizes (or ``graph=False``) run eagerly.  The returned loss is a 0-dim device
    tensor (a static buffer in graph mode: read it before the next call)."""

    def __init__(self, models, embeddings, state: FlatTrainState, loss_fn, *, batch: int, N_samples=64,
                 N_importance=64, use_disp=False, perturb=1.0, noise_std=0.0, white_back=True, chunk=32 * 1024,
                 graph=True, warmup=3, fuse_loss=True, split_adam=True):
        self.models, self.embeddings, self.state, self.loss_fn = models, embeddings, state, loss_fn
        self.batch = int(batch)
        self.args = (int(N_samples), bool(use_disp), float(perturb), float(noise_std), int(N_importance), int(chunk),
                     bool(white_back))
        self.dev = state.flat_p.device
        self.graph = None
        self.launches_per_step = None
        self._warmup = warmup
        self._want_graph = bool(graph) and self.dev.type == "cuda"
        self.terms = None            # loss terms of the last call (dict of 0-dim tensors)
        from .objective import NerfWLoss
        self._fused_loss = bool(fuse_loss) and type(loss_fn) is NerfWLoss
        self._split_adam = bool(split_adam) and dist_info()[1] == 1
        self._unit = torch.ones((), device=self.dev, dtype=torch.float32)      # the standing cotangent of the loss

    # ---- the iteration itself ---------------------------------------------------------------------
    def _iteration(self, rays, ts, rgbs, graph_safe_rng: bool, do_step: bool = True):
        Ns, use_disp, perturb, noise_std, Ni, chunk, white_back = self.args
        kw = {}
        if graph_safe_rng:
            n = rays.shape[0]
            rng = {}
            if perturb > 0:
                rng["perturb_rand"] = torch.rand(n, Ns, device=rays.device)
                if Ni > 0:
                    rng["u"] = torch.rand(n, Ni, device=rays.device)
            kw["rng"] = rng
        if do_step and graph_safe_rng and self._split_adam:
            # one GPU: the fine model (its backward chain is started first and finishes first, [B200] tools/graph_timeline.py)
            # is updated as soon as its weight gradients are done, next to the coarse chain's last kernels; the step's main
            # optimizer launch then covers the coarse model, the embeddings and the step state
            kw["after_fine_chain"] = lambda: self.state.adam_step_early([self.models["fine"]])
        if self._fused_loss:
            # NerfWLoss inside the render node: its backward is formed in the compositing kernels (rendering.render_rays)
            kw["loss"] = dict(fn=self.loss_fn, targets=rgbs, join_in_backward=True)
        res = render_rays(self.models, self.embeddings, rays, ts, Ns, use_disp, perturb, noise_std, Ni, chunk,
                          white_back, **kw)
        terms = res["loss_terms"] if self._fused_loss else self.loss_fn(res, rgbs)
        loss = getattr(terms, "total", None)
        if loss is None:
            loss = sum(terms.values())               # train.py:166
        loss.backward(gradient=self._unit)           # a standing cotangent of 1: no fill launch per step
        if do_step:
            self.state.step()
        return loss.detach(), {k: v.detach() for k, v in terms.items()}

    def _capture(self, rays, ts, rgbs):
        from . import kernels as K
        self.s_rays = rays.clone()
        self.s_ts = ts.clone()
        self.s_rgbs = rgbs.clone()
        # One process per GPU with world > 1: the gradient all-reduce stays OUTSIDE the graph.  The captured part is
        # forward + loss + backward (everything up to the flat gradient buffer); the NCCL all-reduce and the two
        # Adam kernels are enqueued behind the replay on the same stream (3 host calls per iteration instead of
        # ~60).  A collective inside a capture ties the graph to the communicator's internal streams and to the
        # process group's watchdog thread, and a hang there takes the whole job down.
        # With FlatTrainState.enable_p2p() the exchange is one kernel of ours over peer memory (csrc/p2p.cu): no NCCL call, so
        # the whole iteration -- all-reduce and Adam included -- is captured.
        self._split = dist_info()[1] > 1 and not getattr(self.state, "p2p_enabled", False)
        # warm-up iterations (allocator pools, lazy module state) run on a side stream as torch's capture recipe asks;
        # they must not train: one call = one optimizer update (train.py:158-174), so the model, the Adam moments and
        # the device-resident step counter are put back afterwards.  do_step=False also keeps a warm-up from issuing
        # collectives other ranks do not match.
        snap = self.state.snapshot() if self._warmup > 0 else None
        side = torch.cuda.Stream(device=self.dev)
        side.wait_stream(torch.cuda.current_stream(self.dev))
        with torch.cuda.stream(side):
            for _ in range(self._warmup):
                self._iteration(self.s_rays, self.s_ts, self.s_rgbs, True, do_step=False)
        torch.cuda.current_stream(self.dev).wait_stream(side)
        torch.cuda.synchronize(self.dev)
        if snap is not None:
            self.state.restore(snap)
        if self.state.refreshes_caches:
            self.state.prepare()                       # pointer tables + rest list (synchronises): not inside a capture
        else:
            for m in self.models.values():
                m._nwb_dirty = True                    # the captured iteration must contain the operand-cache refresh
        g = torch.cuda.CUDAGraph()
        l0 = K.launch_count
        # thread_local: other threads of the process (NCCL watchdog, clock sampler) may touch CUDA during the capture
        with torch.cuda.graph(g, capture_error_mode="thread_local"):
            self.s_loss, self.s_terms = self._iteration(self.s_rays, self.s_ts, self.s_rgbs, True, do_step=not self._split)
        self.launches_per_step = K.launch_count - l0 + ((1 if self.state.refreshes_caches else 2) if self._split else 0)
        self.graph = g

    def __call__(self, rays, ts, rgbs):
        # Host batches (pinned memory, as a DataLoader with pin_memory=True hands them over) are accepted as they are: in
        # graph mode each goes to its static device buffer in ONE asynchronous host-to-device copy, with no intermediate
        # device tensor and no device-to-device copy behind it.
        if self._want_graph and rays.shape[0] == self.batch:
            if self.graph is None:
                self._capture(*(t.to(self.dev, non_blocking=True) for t in (rays, ts, rgbs)))   # warm-up + capture (the capture itself does not run)
            self.state.pull_lr()                       # a torch scheduler may have moved FlatAdam's lr between steps
            if rays.is_cuda and ts.is_cuda and rgbs.is_cuda and ts.dtype == self.s_ts.dtype and rays.dtype == rgbs.dtype == torch.float32 \
                    and rays.is_contiguous() and ts.is_contiguous() and rgbs.is_contiguous():
                # a resident batch: one launch instead of three copies
                K.copy_segments(((self.s_rays, rays), (self.s_ts, ts), (self.s_rgbs, rgbs)))
            else:                                      # host batches (pinned): three asynchronous host-to-device copies
                self.s_rays.copy_(rays, non_blocking=True)
                self.s_ts.copy_(ts, non_blocking=True)
                self.s_rgbs.copy_(rgbs, non_blocking=True)
            self.graph.replay()
            if self._split:
                self.state.step()                      # all-reduce + Adam, stream-ordered behind the replay
            else:
                self.state.step_count += 1
            if not self.state.refreshes_caches:
                for m in self.models.values():
                    m._nwb_dirty = True                # parameters moved: eager callers must repack
            self.terms = self.s_terms
            return self.s_loss
        rays, ts, rgbs = (t.to(self.dev, non_blocking=True) for t in (rays, ts, rgbs))
        loss, self.terms = self._iteration(rays, ts, rgbs, False)
        return loss
