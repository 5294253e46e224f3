"""Host-side data-parallel logic on CPU: ray-block sharding, the gather of per-ray results and the
flat-buffer gradient all-reduce + Adam, at world_size 2 over gloo; plus the objective
module against the oracle's restatement of losses.py."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from nerf_fl_b200 import objective
from nerf_fl_b200.parallel import FlatAdam, FlatTrainState, gather_rays, shard_range
from oracle import nerfw_oracle as O
from tests.host_models import adam_update_torch, color_loss, nerfw_terms


def test_shard_range_partitions_exactly():
    for n in (0, 1, 7, 640000, 640001):
        for world in (1, 2, 3, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1


def _toy_modules(seed):
    torch.manual_seed(seed)
    return [torch.nn.Linear(5, 3), torch.nn.Embedding(7, 4)]


def test_flat_state_adam_matches_torch_adam_single_process():
    mods, ref = _toy_modules(0), _toy_modules(0)
    state = FlatTrainState(mods, lr=5e-4, host_update=adam_update_torch)
    opt = torch.optim.Adam([p for m in ref for p in m.parameters()], lr=5e-4, eps=1e-8)
    for step in range(3):
        x = torch.randn(6, 5, generator=torch.Generator().manual_seed(step))
        idx = torch.tensor([1, 1, 4])
        for ms, flat in ((mods, True), (ref, False)):
            loss = ms[0](x).pow(2).sum() + ms[1](idx).sum() * 0.3
            grads = torch.autograd.grad(loss, [p for m in ms for p in m.parameters()])
            for p, g in zip([p for m in ms for p in m.parameters()], grads):
                if flat:
                    p.grad.add_(g)                         # kernels accumulate into the flat views
                else:
                    p.grad = g
        state.step()
        opt.step()
        assert state.flat_g.abs().max().item() == 0.0      # zero_grad folded into the step
        for a, b in zip([p for m in mods for p in m.parameters()], [p for m in ref for p in m.parameters()]):
            assert torch.allclose(a, b, atol=1e-7, rtol=1e-6)
    # state_dict keys and shapes are untouched by the re-homing
    assert list(mods[0].state_dict()) == ["weight", "bias"] and mods[1].weight.shape == (7, 4)


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mods = _toy_modules(0)
        state = FlatTrainState(mods, lr=1e-2, host_update=adam_update_torch)
        # each rank contributes its own gradient; DDP semantics = mean over ranks
        for p in state.params:
            p.grad.add_(torch.full_like(p, float(rank + 1)))
        state.step()
        # ray-sharded render: rank r owns a contiguous block, results gathered everywhere
        n = 11
        lo, hi = shard_range(n, rank, world)
        local = torch.arange(lo, hi, dtype=torch.float32)[:, None].repeat(1, 4)
        full = gather_rays(local, n)
        # plain arrays, not torch tensors: a tensor travels through a Queue as a file descriptor that the parent has to fetch
        # from this process while it is still alive (torch.multiprocessing's resource sharer) -- a race once the worker exits
        out.put((rank, state.flat_p.detach().clone().numpy(), full.numpy()))
    finally:
        dist.destroy_process_group()


def test_allreduce_adam_and_gather_world2_gloo():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in procs], key=lambda t: t[0])
    res = [(r, torch.from_numpy(p_), torch.from_numpy(f_)) for r, p_, f_ in res]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # both ranks hold identical parameters after the step
    assert torch.equal(res[0][1], res[1][1])
    # and they equal a single-process Adam step on the mean gradient (1+2)/2
    mods = _toy_modules(0)
    ref = FlatTrainState(mods, lr=1e-2, host_update=adam_update_torch)
    for p in ref.params:
        p.grad.add_(torch.full_like(p, 1.5))
    ref.step()
    assert torch.allclose(res[0][1], ref.flat_p, atol=1e-7)
    want = torch.arange(11, dtype=torch.float32)[:, None].repeat(1, 4)
    assert torch.equal(res[0][2], want) and torch.equal(res[1][2], want)


@pytest.mark.parametrize("with_beta", [True, False])
def test_torch_op_loss_restatement_matches_oracle(with_beta):
    """tests/host_models.nerfw_terms (the autograd reference of the loss kernels' GPU tests) against the oracle's
    restatement of losses.py:35-50 -- and the product refuses CPU tensors."""
    gen = torch.Generator().manual_seed(4)
    res = {"rgb_coarse": torch.rand(40, 3, generator=gen), "rgb_fine": torch.rand(40, 3, generator=gen)}
    if with_beta:
        res["beta"] = torch.rand(40, generator=gen) + 0.1
        res["transient_sigmas"] = torch.rand(40, 128, generator=gen)
    tgt = torch.rand(40, 3, generator=gen)
    got = nerfw_terms(res, tgt)
    want = O.nerfw_loss(res, tgt)
    assert set(got) == set(want)
    for k in want:
        assert torch.allclose(got[k], want[k], atol=0, rtol=1e-6), k
    c = color_loss(res, tgt)
    assert torch.allclose(c, ((res["rgb_coarse"] - tgt) ** 2).mean() + ((res["rgb_fine"] - tgt) ** 2).mean())
    with pytest.raises(RuntimeError, match="no CPU path"):
        objective.NerfWLoss()(res, tgt)
    with pytest.raises(RuntimeError, match="no CPU path"):
        objective.ColorLoss()(res, tgt)


def test_flat_state_without_host_update_refuses_cpu():
    state = FlatTrainState(_toy_modules(0), lr=1e-3)
    with pytest.raises(RuntimeError, match="no CPU"):
        state.step()


def test_flat_adam_state_dict_round_trip():
    """Resume (train.py:231): the Adam moments and the step live in the flat buffers; FlatAdam.state_dict exports
    them in torch.optim.Adam's layout and load_state_dict restores them -- a resumed run continues the trajectory."""
    def grads(state, k):
        g = torch.Generator().manual_seed(k)
        for p in state.params:
            p.grad.copy_(torch.randn(p.shape, generator=g))

    a = FlatTrainState(_toy_modules(0), lr=1e-2, host_update=adam_update_torch)
    opt_a = FlatAdam(a)
    for k in range(3):
        grads(a, k)
        opt_a.step()
    ckpt = {"opt": opt_a.state_dict(), "model": [m.state_dict() for m in a.modules]}
    assert set(ckpt["opt"]["state"][0]) == {"step", "exp_avg", "exp_avg_sq"} and float(ckpt["opt"]["state"][0]["step"]) == 3
    # torch.optim.Adam accepts the exported state
    ref_mods = _toy_modules(0)
    torch.optim.Adam([p for m in ref_mods for p in m.parameters()], lr=1e-2).load_state_dict(
        {"state": ckpt["opt"]["state"], "param_groups": ckpt["opt"]["param_groups"]})
    mods_b = _toy_modules(1)
    for m, sd in zip(mods_b, ckpt["model"]):
        m.load_state_dict(sd)
    b = FlatTrainState(mods_b, lr=5e-4, host_update=adam_update_torch)
    opt_b = FlatAdam(b)
    opt_b.load_state_dict(ckpt["opt"])
    assert b.step_count == 3 and abs(b.lr - 1e-2) < 1e-12
    for k in range(3, 5):
        grads(a, k); opt_a.step()
        grads(b, k); opt_b.step()
    assert torch.equal(a.flat_p, b.flat_p) and torch.equal(a.exp_avg_sq, b.exp_avg_sq)
