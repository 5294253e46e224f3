"""Print selected per-kernel rows of a bench.py JSON line read from stdin: python tools/bench_kern.py composite sample"""
import json, sys
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print("value", round(d["value"]), "ms", round(d["ms_per_step"], 4))
for k in d["kernels"]:
    if any(p in k["kernel"] for p in sys.argv[1:]):
        print(" ", k["kernel"], k["mode"], "avg_ms", k["avg_ms"], "frac", k.get("frac"), k.get("achieved"), k.get("unit"))
