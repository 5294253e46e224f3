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
