"""nwb_nerfw_loss_fwd_static -- the loss forward with its scratch in a static device array of the library (no counterpart
in the reference).  Collected last on purpose (pytest -x)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_RAYS = 1 << 20          # NWB_LOSS_STATIC_MAX_RAYS, include/nerfw_b200.h


def _bind():
    from nerf_fl_b200 import _lib as L
    vp = C.c_void_p
    return L, L.bind("nwb_nerfw_loss_fwd_static", [vp] * 5 + [C.c_int, C.c_int64, C.c_float, C.c_float, vp, vp], C.c_int)


def _inputs(n, S, seed, dev):
    g = torch.Generator().manual_seed(seed)
    rgb_c, rgb_f, tgt = (torch.rand(n, 3, generator=g).to(dev) for _ in range(3))
    beta = (torch.rand(n, generator=g) + 0.1).to(dev)
    ts = torch.rand(n, S, generator=g).to(dev)
    return rgb_c, rgb_f, beta, ts, tgt


@pytest.mark.parametrize("n,S", [(1024, 256), (37, 5), (40000, 8)])
def test_static_scratch_loss_matches_caller_scratch_loss_twice_in_a_row(n, S):
    from nerf_fl_b200 import kernels as K
    L, fn = _bind()
    dev = torch.device("cuda:0")
    for call in range(2):                                   # the second call finds the first call's contents in the array
        rgb_c, rgb_f, beta, ts, tgt = _inputs(n, S, 100 + call, dev)
        want = K.nerfw_loss_fwd(rgb_c, rgb_f, beta, ts, tgt, 0.01, 1.0)
        terms = torch.full((5,), float("nan"), device=dev)
        L.check(fn(L.ptr(rgb_c), L.ptr(rgb_f), L.ptr(beta), L.ptr(ts), L.ptr(tgt), n, ts.numel(), 0.01, 1.0,
                   L.ptr(terms), L.stream_ptr(dev)), "nwb_nerfw_loss_fwd_static")
        torch.cuda.synchronize()
        assert torch.equal(terms, want), (call, terms, want)          # same kernel, same grid, same order: bit-identical
        # variants without the transient head / without the fine pass
        for f, b, t in ((rgb_f, None, None), (None, None, None)):
            want = K.nerfw_loss_fwd(rgb_c, f, b, t, tgt, 0.01, 1.0)
            L.check(fn(L.ptr(rgb_c), L.ptr(f), L.ptr(b), L.ptr(t), L.ptr(tgt), n, 0, 0.01, 1.0, L.ptr(terms),
                       L.stream_ptr(dev)), "nwb_nerfw_loss_fwd_static")
            assert torch.equal(terms, want)


def test_static_scratch_loss_rejects_sizes_beyond_its_maximum():
    L, fn = _bind()
    dev = torch.device("cuda:0")
    x = torch.zeros(8, 3, device=dev)
    terms = torch.zeros(5, device=dev)
    rc = fn(L.ptr(x), None, None, None, L.ptr(x), MAX_RAYS + 1, 0, 0.0, 1.0, L.ptr(terms), L.stream_ptr(dev))
    assert rc != 0
    with pytest.raises(RuntimeError, match="maximum"):
        L.check(rc, "nwb_nerfw_loss_fwd_static")
    torch.cuda.synchronize()
