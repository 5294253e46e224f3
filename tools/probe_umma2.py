"""Interference probe: cycles per tcgen05.mma (M=128 N=256 K=16, floor 128) while other warps write shared memory,
fence, read TMEM, or stream bulk copies from L2.  mode bits: 1 st.shared, 2 fence.proxy.async, 4 bulk g2s stream,
8 tcgen05.ld, 16 paced, 32 / 64: each 32 KB stage as 8 x 4 KB / 2 x 16 KB copies."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_fl_b200 import _lib as L
from tools.build_probes import build
_P = C.CDLL(build())          # tools/libnwb_probes.so (not part of the product library)
fn = _P.nwb_debug_umma_probe2
fn.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]; fn.restype = C.c_int
src = torch.zeros(40 * 32768, dtype=torch.uint8, device="cuda")
rounds = 256
for grid in (1, 148):
    out = torch.zeros(2 * grid, dtype=torch.int64, device="cuda")
    for mode in (0, 1, 3, 1 | 16, 3 | 16, 8, 8 | 16, 4, 4 | 64, 4 | 32, 4 | 1, 4 | 3 | 8, 4 | 3 | 8 | 16, 4 | 3 | 8 | 16 | 32):
        for rep in range(2):
            out.zero_()
            assert 0 == (fn(grid, rounds, mode, src.data_ptr(), out.data_ptr(), torch.cuda.current_stream().cuda_stream))
            torch.cuda.synchronize()
        cyc = out[:grid].float().mean().item()
        byt = out[grid:].float().mean().item()
        print(f"grid={grid:3d} mode={mode:2d}: {cyc / (rounds * 4):6.1f} cyc/mma   bulk ingest {byt / max(cyc, 1):5.1f} B/cyc/SM")
