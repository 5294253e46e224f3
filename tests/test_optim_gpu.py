"""(f2) Adam on the flat buffers: the launch that also refreshes the packed operand caches
(nwb_adam_pack_step_dev) against nwb_adam_step_dev followed by nwb_field_pack, and both against
torch.optim.Adam (the optimizer utils/__init__.py:24-33 builds for the reference)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

if not torch.cuda.is_available():
    pytest.skip("needs a GPU", allow_module_level=True)

from nerf_fl_b200 import synthetic as syn                      # noqa: E402
from nerf_fl_b200.nerf import NeRF                             # noqa: E402
from nerf_fl_b200.parallel import FlatTrainState               # noqa: E402

DEV = torch.device("cuda:0")


def _modules(seed):
    torch.manual_seed(seed)
    coarse = NeRF("coarse", in_channels_xyz=63, in_channels_dir=27).to(DEV)
    fine = NeRF("fine", in_channels_xyz=63, in_channels_dir=27, encode_appearance=True, in_channels_a=48,
                encode_transient=True, in_channels_t=16, beta_min=0.1).to(DEV)
    emb = [torch.nn.Embedding.from_pretrained(syn.make_embedding(s, 20, d), freeze=False).to(DEV) for s, d in ((3, 48), (4, 16))]
    return [coarse, fine] + emb


def _grads(n, steps, seed):
    g = torch.Generator().manual_seed(seed)
    return [(torch.randn(n, generator=g) * 10.0 ** float(torch.randint(-6, 1, (1,), generator=g))).to(DEV)
            for _ in range(steps)]


def test_fused_adam_pack_equals_adam_then_pack():
    a = FlatTrainState(_modules(7), lr=5e-4, fused_pack=True)
    b = FlatTrainState(_modules(7), lr=5e-4, fused_pack=False)
    assert a.refreshes_caches and not b.refreshes_caches
    assert torch.equal(a.flat_p, b.flat_p)
    for i, g in enumerate(_grads(a.numel, 4, 11)):
        a.flat_g.copy_(g)
        b.flat_g.copy_(g)
        a.adam_step(world=2)
        b.adam_step(world=2)
        torch.cuda.synchronize()
        assert torch.equal(a.flat_p, b.flat_p), f"step {i}: parameters differ"
        assert torch.equal(a.exp_avg, b.exp_avg) and torch.equal(a.exp_avg_sq, b.exp_avg_sq)
        assert float(a.flat_g.abs().max()) == 0.0                     # zero_grad folded into the same pass
        assert torch.equal(a.hyper[:4], b.hyper[:4])                  # device step state advanced identically
        for ma, mb in zip(a.modules[:2], b.modules[:2]):
            assert not ma._nwb_dirty and mb._nwb_dirty
            pa = ma._nwb_cache[1]                                      # written by the optimizer launch
            pb = mb.nwb_packed()                                       # repacked by nwb_field_pack
            assert torch.equal(pa, pb), f"step {i}: packed operand cache differs"
    # every flat offset is updated exactly once: cache-covered + rest = all
    fz = a._fused
    assert fz["covered"] + fz["rest"].numel() == a.numel
    assert fz["covered"] > 1_100_000 and fz["rest"].numel() < 120_000


def test_fused_adam_matches_torch_adam():
    mods = _modules(9)
    st = FlatTrainState(mods, lr=5e-4)
    ref_p = st.flat_p.clone().requires_grad_(True)
    opt = torch.optim.Adam([ref_p], lr=5e-4, eps=1e-8, weight_decay=0.0)
    for g in _grads(st.numel, 5, 13):
        st.flat_g.copy_(g)
        ref_p.grad = g.clone()
        opt.step()
        st.adam_step()
    torch.cuda.synchronize()
    d = (st.flat_p - ref_p.detach()).abs().max().item()
    assert d <= 2e-6, d                                                # fp32 rounding of a 5e-4 step


def test_fused_adam_rejects_foreign_parameters():
    """A field whose parameters are not all in the flat buffer stays on the repack path."""
    mods = _modules(3)
    for p in mods[0].xyz_encoding_1.parameters():
        p.requires_grad_(False)
    st = FlatTrainState(mods, lr=1e-3)
    assert len(st._fused_fields) == 1 and st._fused_fields[0] is mods[1]
    st.flat_g.normal_()
    st.adam_step()
    assert mods[0]._nwb_dirty or mods[0]._nwb_cache is None
    assert not mods[1]._nwb_dirty


def test_optimizer_step_split_over_two_launches_is_bit_identical():
    """adam_step_early (NWB_ADAM_NO_ADVANCE) updates one model ahead of the step's main launch: parameters, moments, packed
    operand caches and the device step state come out exactly as from the single launch."""
    a = FlatTrainState(_modules(7), lr=5e-4)
    b = FlatTrainState(_modules(7), lr=5e-4)
    for i, g in enumerate(_grads(a.numel, 4, 17)):
        a.flat_g.copy_(g)
        b.flat_g.copy_(g)
        a.adam_step()
        assert b.adam_step_early([b.modules[0]])
        assert not b.adam_step_early([b.modules[0]])                   # once per step
        b.adam_step()
        torch.cuda.synchronize()
        assert torch.equal(a.flat_p, b.flat_p), f"step {i}: parameters differ"
        assert torch.equal(a.exp_avg, b.exp_avg) and torch.equal(a.exp_avg_sq, b.exp_avg_sq)
        assert float(b.flat_g.abs().max()) == 0.0
        assert torch.equal(a.hyper, b.hyper)
        for ma, mb in zip(a.modules[:2], b.modules[:2]):
            assert torch.equal(ma._nwb_cache[1], mb._nwb_cache[1]), f"step {i}: packed operand cache differs"
    assert a.step_count == b.step_count == 4
