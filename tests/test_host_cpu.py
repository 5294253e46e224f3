"""CPU-side checks: the C-ABI library loads and exports every symbol include/nerfw_b200.h declares,
and the host-side layout helpers round-trip."""
import ctypes
import os
import re

import torch

from nerf_fl_b200 import layout

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "nerfw_b200.h")).read()
    names = sorted(set(re.findall(r"\b(nwb_[a-z0-9_]+)\s*\(", hdr)))
    assert len(names) >= 8
    lib = ctypes.CDLL(os.path.join(ROOT, "nerf_fl_b200", "libnerfw_b200.so"))
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    lib.nwb_abi_version.restype = ctypes.c_int
    assert lib.nwb_abi_version() == 1


def test_tile_image_roundtrip_and_swizzle():
    m = torch.arange(16 * 128, dtype=torch.float32).reshape(16, 128) % 251
    img = layout.tile_image(m)
    assert torch.equal(layout.from_tile_image(img, 16, 128).float(), m)
    off = layout.image_offsets(16, 128)
    assert sorted(off.flatten().tolist()) == list(range(16 * 128))       # a permutation
    # row 0 is unswizzled; row 1's first chunk lands in chunk slot 1; next column block at R*64
    assert off[0, :64].tolist() == list(range(64))
    assert off[1, 0].item() == 64 + 8 and off[1, 8].item() == 64
    assert off[0, 64].item() == 16 * 64 and off[8, 0].item() == 512


def test_flat_adam_facade_follows_torch_schedulers():
    """utils/__init__.py:44-61: the reference builds MultiStepLR / CosineAnnealingLR / LambdaLR over its optimizer;
    FlatAdam lets the same schedulers drive the flat-buffer Adam.  CPU tensors: the update is the torch-op Adam the tests inject (tests/host_models.py)."""
    import torch
    from torch.optim.lr_scheduler import CosineAnnealingLR, LambdaLR, MultiStepLR
    from nerf_fl_b200.parallel import FlatAdam, FlatTrainState
    from tests.host_models import adam_update_torch

    def run(make_sched):
        torch.manual_seed(0)
        lin_a, lin_b = torch.nn.Linear(5, 7), torch.nn.Linear(5, 7)
        lin_b.load_state_dict(lin_a.state_dict())
        st = FlatTrainState([lin_a], lr=5e-4, host_update=adam_update_torch)
        opt_a, opt_b = FlatAdam(st), torch.optim.Adam(lin_b.parameters(), lr=5e-4, eps=1e-8)
        sch_a, sch_b = make_sched(opt_a), make_sched(opt_b)
        g = torch.Generator().manual_seed(1)
        for epoch in range(6):
            for _ in range(3):
                for pa, pb in zip(lin_a.parameters(), lin_b.parameters()):
                    gr = torch.randn(pa.shape, generator=g)
                    pa.grad.copy_(gr)                       # FlatTrainState's preassigned views
                    pb.grad = gr.clone()
                opt_a.step(); opt_b.step()
                opt_a.zero_grad(); opt_b.zero_grad(set_to_none=False)
                assert lin_a.weight.grad is not None and lin_a.weight.grad.data_ptr() == st.flat_g.data_ptr()
            sch_a.step(); sch_b.step()
            assert abs(opt_a.param_groups[0]["lr"] - opt_b.param_groups[0]["lr"]) < 1e-12
        for pa, pb in zip(lin_a.parameters(), lin_b.parameters()):
            assert (pa - pb).abs().max().item() <= 1e-6
        return st

    run(lambda o: MultiStepLR(o, milestones=[2, 4], gamma=0.1))
    run(lambda o: CosineAnnealingLR(o, T_max=6, eta_min=1e-8))
    st = run(lambda o: LambdaLR(o, lambda epoch: (1 - epoch / 6) ** 0.9))
    st.pull_lr()
    assert st.lr < 5e-4
