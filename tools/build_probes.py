"""Builds tools/libnwb_probes.so: the tcgen05 issue-cost / interference probes (tools/csrc/umma_probe*.cu).
Measurement code only -- kept out of libnerfw_b200.so (the product library exports what include/nerfw_b200.h declares
and nothing else).  Usage: python tools/build_probes.py ; then tools/probe_umma.py / tools/probe_umma2.py."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(ROOT, "nerf_fl_b200", "csrc")
LIB = os.path.join(HERE, "libnwb_probes.so")


def build() -> str:
    srcs = [os.path.join(HERE, "csrc", f) for f in ("umma_probe.cu", "umma_probe2.cu")] + [os.path.join(CSRC, "api.cu")]
    cmd = ["/usr/local/cuda/bin/nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
           "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-I", CSRC, "-shared", "-o", LIB] + srcs
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("probe build failed")
    return LIB


if __name__ == "__main__":
    print(build())
