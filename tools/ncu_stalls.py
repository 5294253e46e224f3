"""Warp-stall sampling summary of one kernel from an `ncu --set full --import-source on` report.
usage: python tools/ncu_stalls.py report.ncu-rep kernel_regex [launch_index] > profiles/<name>.md"""
import csv
import subprocess
import sys

rep, rx = sys.argv[1], sys.argv[2]
skip = sys.argv[3] if len(sys.argv) > 3 else "0"
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{rx}", "--launch-skip", skip,
                      "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
name, h, data = rows[0][1], rows[1], rows[2:]
ci = {n: i for i, n in enumerate(h)}
stalls = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
tot = {n: 0 for n in stalls}
total, recs = 0, []
for r in data:
    if len(r) < len(h) or not (r[ci["# Samples"]] or "0").isdigit():
        continue
    ns = int(r[ci["# Samples"]] or 0)
    total += ns
    st = {n: int(r[ci[n]] or 0) for n in stalls}
    for n in stalls:
        tot[n] += st[n]
    recs.append((ns, r[ci["Source"]].strip(), st))
print(f"# warp stall sampling: `{name}` ({rep}, launch {skip}; {total} samples over all warps of the CTA roles)\n")
print("| stall reason | samples | share |\n|---|---:|---:|")
for n, v in sorted(tot.items(), key=lambda x: -x[1])[:12]:
    print(f"| {n} | {v} | {100 * v / max(total, 1):.1f} % |")
print("\n| samples | share | SASS | top reasons |\n|---:|---:|---|---|")
for ns, src, st in sorted(recs, key=lambda x: -x[0])[:25]:
    top = ", ".join(f"{k[6:]} {v}" for k, v in sorted(st.items(), key=lambda x: -x[1])[:2] if v)
    print(f"| {ns} | {100 * ns / max(total, 1):.1f} % | `{src[:80]}` | {top} |")
