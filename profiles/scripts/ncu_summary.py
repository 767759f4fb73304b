"""Summarise an `ncu --set full` report (read with `ncu -i X.ncu-rep --page raw --csv > raw.csv`) as a markdown table.
Usage: python profiles/scripts/ncu_summary.py raw.csv > profiles/rNN_x_ncu_kernels.md"""
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, data = rows[0], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
cols = [("gpu__time_duration.sum", "us"), ("dram__bytes_read.sum", "rd MB"), ("dram__bytes_write.sum", "wr MB"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor %"),
        ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "fp64 %"),
        ("sm__inst_issued.avg.pct_of_peak_sustained_active", "issue %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps %"),
        ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid"), ("launch__block_size", "block")]
print("| # | kernel | " + " | ".join(c[1] for c in cols) + " |")
print("|---|---|" + "---|" * len(cols))
for n, r in enumerate(data):
    name = re.sub(r"\(.*", "", r[ix["Kernel Name"]]).replace("void ", "")
    vals = []
    for k, _ in cols:
        v = r[ix[k]] if k in ix else ""
        try:
            v = f"{float(v):.1f}" if "." in v else v
        except ValueError:
            pass
        vals.append(v)
    print(f"| {n} | `{name}` | " + " | ".join(vals) + " |")
