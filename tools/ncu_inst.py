"""Executed warp-instruction counts of one kernel from an `ncu --set full --import-source on` report: total, per
opcode, and the cycles they imply (4 schedulers per SM).
usage: python tools/ncu_inst.py report.ncu-rep kernel_regex [launch_index]"""
import collections, csv, subprocess, sys
rep, rx = sys.argv[1], sys.argv[2]
skip = sys.argv[3] if len(sys.argv) > 3 else "0"
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{rx}", "--launch-skip", skip,
                      "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, data = rows[1], rows[2:]
ci = {n: i for i, n in enumerate(h)}
tot = 0
by = collections.Counter()
for r in data:
    if len(r) < len(h) or not (r[ci["# Samples"]] or "0").isdigit():
        continue
    n = int(r[ci["Instructions Executed"]] or 0)
    tot += n
    src = r[ci["Source"]].strip().split()
    op = src[1] if src and src[0].startswith("@") else (src[0] if src else "?")
    by[op.split(".")[0]] += n
print(f"{rows[0][1]}: {tot} warp-instructions executed")
for op, n in by.most_common(40):
    print(f"  {op:12s} {n:10d}  {100 * n / tot:5.1f} %")
