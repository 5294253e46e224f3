"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel:
launches, total/avg device time and share (cold-cache, serialised: compare SHARES, not absolutes).
usage: python tools/summarize_launches.py gpurun_out/launches.csv [skip_first_n] > profiles/<name>.md"""
import collections
import csv
import re
import sys

path = sys.argv[1]
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
rows = [r for r in csv.reader(open(path)) if len(r) > 14 and r[0].isdigit()]
rows = rows[skip:]
agg = collections.OrderedDict()
for r in rows:
    name = re.sub(r"\(.*", "", r[4]).replace("void ", "")
    name = re.sub(r"<.*", "<...>", name)
    ours = not name.startswith(("at::", "nccl", "cub::")) and "at::native" not in r[4]
    key = (name, r[7], r[8])
    n, ns, _ = agg.get(key, (0, 0.0, ours))
    agg[key] = (n + 1, ns + float(r[14]), ours)
total = sum(v[1] for v in agg.values())
print(f"# launch list summary of `{path}` ({len(rows)} launches, first {skip} skipped)\n")
print("| kernel | block | grid | launches | total us | avg us | share | ours |")
print("|---|---|---|---:|---:|---:|---:|---|")
for (name, blk, grid), (n, ns, ours) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"| `{name}` | {blk} | {grid} | {n} | {ns / 1e3:.1f} | {ns / 1e3 / n:.2f} | {ns / total * 100:.1f}% | {'yes' if ours else 'torch'} |")
mine = sum(v[1] for v in agg.values() if v[2])
print(f"\nown kernels: {mine / total * 100:.1f}% of device time; torch (loss, allocator fills, optimizer glue): {(1 - mine / total) * 100:.1f}%")
