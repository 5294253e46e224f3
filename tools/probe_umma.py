"""Issue-path probe: cycles for one thread to issue tcgen05.mma (M=128, N, K=16) and for them to drain."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_fl_b200 import _lib as L
from tools.build_probes import build
_P = C.CDLL(build())          # tools/libnwb_probes.so (not part of the product library)
fn = _P.nwb_debug_umma_probe
fn.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]; fn.restype = C.c_int
out = torch.zeros(2, dtype=torch.int64, device="cuda")
for N in (64, 128, 256):
    for commit_every in (0, 1):
        for rounds in (4, 64):
            for rep in range(2):
                assert 0 == (fn(N, rounds, commit_every, out.data_ptr(), torch.cuda.current_stream().cuda_stream))
                torch.cuda.synchronize()
            issue, total = out.tolist()
            n = rounds * 4
            print(f"N={N:3d} commit_every={commit_every} mmas={n:3d}: issue {issue:6d} cyc ({issue / n:6.1f}/mma)  drained {total:6d} cyc ({total / n:6.1f}/mma, floor {N // 2})")
