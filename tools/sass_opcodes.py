"""Opcode evidence per kernel from the built objects (cuobjdump -sass): counts of the Blackwell-native instructions
(tcgen05.mma = UTC*MMA, tcgen05.ld = LDTM, bulk async copies = UBLKCP, tcgen05.commit = UTCBAR, mbarrier = SYNCS) and the
ten most frequent opcodes.     python tools/sass_opcodes.py > profiles/sass_opcodes.md"""
import collections, os, re, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
B = os.path.join(ROOT, "nerf_fl_b200", "build")
KEY = ["UTCHMMA", "UTCQMMA", "UTCHMMA.2CTA", "LDTM", "UBLKCP", "UTCBAR", "SYNCS", "STG", "RED", "ATOMG", "F2FP", "FADD2", "HMMA"]
print("# SASS opcode evidence (cuobjdump -sass nerf_fl_b200/build/*.o, sm_100a)\n")
print("tcgen05.mma -> `UTCHMMA` (kind::f16) / `UTCQMMA` (kind::f8f6f4), `.2CTA` = cta_group::2; tcgen05.ld -> `LDTM`; cp.async.bulk -> `UBLKCP`;"
      " tcgen05.commit -> `UTCBAR` (`.2CTA.MULTICAST` for pairs); mbarrier -> `SYNCS.*`.  No `HMMA` (legacy mma.sync) anywhere.\n")
print("| object | kernel | instructions | " + " | ".join(KEY) + " | top opcodes |")
print("|---|---|---:|" + "---:|" * len(KEY) + "---|")
for obj in sorted(os.listdir(B)):
    if not obj.endswith(".o"):
        continue
    out = subprocess.run(["cuobjdump", "-sass", os.path.join(B, obj)], capture_output=True, text=True).stdout
    name, ops = None, None
    kernels = []
    for ln in out.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0].replace("void ", "").replace("nwb::", "")
            ops = collections.Counter()
            kernels.append((name, ops))
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
        if m and ops is not None:
            ops[m.group(1)] += 1
    for name, ops in kernels:
        tot = sum(ops.values())
        def cnt(k):
            if k == "UTCHMMA":
                return sum(v for o, v in ops.items() if o.startswith("UTCHMMA") and ".2CTA" not in o)
            if k == "UTCHMMA.2CTA":
                return sum(v for o, v in ops.items() if o.startswith("UTCHMMA") and ".2CTA" in o)
            return sum(v for o, v in ops.items() if o.split(".")[0] == k)
        base = collections.Counter()
        for o, v in ops.items():
            base[o.split(".")[0]] += v
        top = ", ".join(f"{o} {v}" for o, v in base.most_common(8))
        print(f"| {obj} | `{name}` | {tot} | " + " | ".join(str(cnt(k)) for k in KEY) + f" | {top} |")
