"""Program-order listing of the SASS lines around the hottest stall sites of one kernel in an ncu report.
usage: python tools/ncu_hot.py report.ncu-rep kernel_regex [launch_index] [min_samples] [context]"""
import csv, subprocess, sys
rep, rx = sys.argv[1], sys.argv[2]
skip = sys.argv[3] if len(sys.argv) > 3 else "0"
thr = int(sys.argv[4]) if len(sys.argv) > 4 else 150
ctx = int(sys.argv[5]) if len(sys.argv) > 5 else 6
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{rx}", "--launch-skip", skip,
                      "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, data = rows[1], rows[2:]
ci = {n: i for i, n in enumerate(h)}
stalls = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
hot = [i for i, r in enumerate(data) if len(r) >= len(h) and (r[ci["# Samples"]] or "0").isdigit() and int(r[ci["# Samples"]] or 0) >= thr]
shown = set()
for i in hot:
    for j in range(max(0, i - ctx), min(len(data), i + 2)):
        if j in shown:
            continue
        shown.add(j)
        r = data[j]
        if len(r) < len(h):
            continue
        ns = int(r[ci["# Samples"]] or 0)
        st = {n: int(r[ci[n]] or 0) for n in stalls}
        top = ", ".join(f"{k[6:]} {v}" for k, v in sorted(st.items(), key=lambda x: -x[1])[:2] if v)
        print(f"{j:6d} {ns:6d}  {r[ci['Source']].strip()[:90]:90s} {top}")
    print("   ...")
