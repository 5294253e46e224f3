"""Markdown table of the metrics that matter from an `ncu --set full` report.
usage: python tools/ncu_table.py gpurun_out/prof.ncu-rep "title" > profiles/<name>.md"""
import csv
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes.sum.per_second",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "sm__cycles_elapsed.max"]
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
cols = [hdr.index(w) for w in WANT if w in hdr]
kn = hdr.index("Kernel Name")
print(f"# {sys.argv[2] if len(sys.argv) > 2 else sys.argv[1]}\n")
print("| kernel | " + " | ".join(f"{hdr[c]} [{units[c]}]" for c in cols) + " |")
print("|---|" + "---|" * len(cols))
for r in rows[2:]:
    name = r[kn].split("(")[0].replace("void ", "").replace("nwb::", "")
    print(f"| `{name}` | " + " | ".join(r[c] for c in cols) + " |")
