"""Host-side description of the bf16 "tile image" layout shared by every tcgen05 operand.

A tile image of an (R x Cc) matrix (R % 8 == 0, Cc % 64 == 0) is the byte image UMMA's
SWIZZLE_128B canonical layout expects in shared memory, so a linear bulk copy lands it ready to
use: column block cb (64 columns = 128 B per row) starts at cb*R*128 bytes; inside a block row r
sits at (r//8)*1024 + (r%8)*128 and its 16-byte chunk j is stored at chunk position j ^ (r%8).
Read K-major (rows = M/N index) or MN-major (rows = contraction index) from the same bytes.
"""
from __future__ import annotations

import torch


def image_offsets(R: int, Cc: int) -> torch.Tensor:
    """(R, Cc) int64: bf16-element offset of element (r, c) inside the tile image."""
    assert R % 8 == 0 and Cc % 64 == 0, (R, Cc)
    r = torch.arange(R).unsqueeze(1)
    c = torch.arange(Cc).unsqueeze(0)
    cb, j, e = c // 64, (c % 64) // 8, c % 8
    return cb * (R * 64) + (r // 8) * 512 + (r % 8) * 64 + ((j ^ (r % 8)) * 8) + e


def tile_image(m: torch.Tensor) -> torch.Tensor:
    """(R, Cc) float tensor -> flat bf16 tensor holding its tile image (same device)."""
    R, Cc = m.shape
    off = image_offsets(R, Cc).to(m.device).flatten()
    out = torch.empty(R * Cc, dtype=torch.bfloat16, device=m.device)
    out[off] = m.to(torch.bfloat16).flatten()
    return out


def from_tile_image(img: torch.Tensor, R: int, Cc: int) -> torch.Tensor:
    """Inverse of tile_image -> (R, Cc) bf16."""
    off = image_offsets(R, Cc).to(img.device).flatten()
    return img.flatten()[off].reshape(R, Cc)


# ---- 8-bit images (e4m3 / e5m2): same swizzle, 128 columns per 128-byte row --------------------------------------
def image8_offsets(R: int, Cc: int) -> torch.Tensor:
    """(R, Cc) int64: BYTE offset of element (r, c) inside an 8-bit tile image (R % 8 == 0, Cc % 128 == 0):
    column block cb (128 columns = 128 B per row) at cb*R*128; row r at (r//8)*1024 + (r%8)*128; 16-byte chunk j
    (16 columns) at chunk position j ^ (r%8)."""
    assert R % 8 == 0 and Cc % 128 == 0, (R, Cc)
    r = torch.arange(R).unsqueeze(1)
    c = torch.arange(Cc).unsqueeze(0)
    cb, j, e = c // 128, (c % 128) // 16, c % 16
    return cb * (R * 128) + (r // 8) * 1024 + (r % 8) * 128 + ((j ^ (r % 8)) * 16) + e


def tile_image8(m: torch.Tensor, dtype=torch.float8_e4m3fn) -> torch.Tensor:
    """(R, Cc) float tensor -> flat uint8 tensor holding its 8-bit tile image (rounded to `dtype`)."""
    R, Cc = m.shape
    off = image8_offsets(R, Cc).to(m.device).flatten()
    out = torch.empty(R * Cc, dtype=torch.uint8, device=m.device)
    out[off] = m.to(dtype).view(torch.uint8).flatten()
    return out


def from_tile_image8(img: torch.Tensor, R: int, Cc: int, dtype=torch.float8_e4m3fn) -> torch.Tensor:
    """Inverse of tile_image8 -> (R, Cc) float32."""
    off = image8_offsets(R, Cc).to(img.device).flatten()
    return img.flatten()[off].view(dtype).float().reshape(R, Cc)


# ---- SWIZZLE_32B 8-bit images: what the field kernels save for the weight-gradient contraction (csrc/field.cuh) ------
def image8s32_offsets(R: int, Cc: int) -> torch.Tensor:
    """SWIZZLE_32B 8-bit image: 32-column group m = c // 32 is (R rows x 32 B) contiguous, row r at r * 32, its two 16-byte
    chunks swapped when bit 2 of r is set (Swizzle<1,4,3>: address bit 4 ^= address bit 7)."""
    assert R % 8 == 0 and Cc % 32 == 0, (R, Cc)
    r = torch.arange(R).unsqueeze(1)
    c = torch.arange(Cc).unsqueeze(0)
    return (c // 32) * (R * 32) + r * 32 + ((((c % 32) // 16) ^ ((r >> 2) & 1)) * 16) + c % 16


def tile_image8s32(m: torch.Tensor, dtype=torch.float8_e4m3fn) -> torch.Tensor:
    R, Cc = m.shape
    off = image8s32_offsets(R, Cc).to(m.device).flatten()
    out = torch.empty(R * Cc, dtype=torch.uint8, device=m.device)
    out[off] = m.to(dtype).view(torch.uint8).flatten()
    return out
