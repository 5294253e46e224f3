"""Build libnerfw_b200.so with extra nvcc flags into tools/variants/<name>.so (objects under tools/variants/<name>/),
for A/B measurements of compile-time kernel options in one gpurun call:
    python tools/build_variant.py sleepy -DNWB_FWD_SLEEPY
    NWB_LIB_PATH=tools/variants/sleepy.so python tools/perf_field.py"""
import os, subprocess, sys
from concurrent.futures import ThreadPoolExecutor
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from nerf_fl_b200 import build as B

name, flags = sys.argv[1], sys.argv[2:]
out = os.path.join(ROOT, "tools", "variants")
objdir = os.path.join(out, name)
os.makedirs(objdir, exist_ok=True)

def one(src):
    obj = os.path.join(objdir, src[:-3] + ".o")
    r = subprocess.run([B.NVCC] + B.FLAGS + flags + ["-c", os.path.join(B.CSRC, src), "-o", obj], capture_output=True, text=True)
    if r.returncode:
        sys.stderr.write(r.stderr); raise SystemExit(1)
    return obj

with ThreadPoolExecutor(8) as ex:
    objs = list(ex.map(one, B._sources()))
lib = os.path.join(out, name + ".so")
subprocess.check_call([B.NVCC, "-shared", "-o", lib] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"])
print(lib)
