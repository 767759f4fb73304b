"""Top stalled SASS instructions of the n-th kernel in an `ncu -i X.ncu-rep --page source --csv` export.
Usage: python profiles/scripts/ncu_source_top.py src.csv [kernel_index] [top_n]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 30
blocks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "hdr": None, "data": []}
        blocks.append(cur)
    elif cur is not None and cur["hdr"] is None:
        cur["hdr"] = r
    elif cur is not None and r and r[0].startswith("0x"):
        cur["data"].append(r)
print(len(blocks), "kernels:", [b["name"] for b in blocks][:8])
b = blocks[which]
ix = {h: i for i, h in enumerate(b["hdr"])}
data = b["data"]
tot = sum(int(r[ix["# Samples"]]) for r in data)
print("kernel", which, b["name"], "instructions", len(data), "samples", tot)
sc = ["stall_long_sb", "stall_short_sb", "stall_wait", "stall_mio", "stall_lg", "stall_membar", "stall_math", "stall_barrier", "stall_sleep",
      "stall_not_selected", "stall_selected"]
sc = [s for s in sc if s in ix]
print("  idx samples     exec " + " ".join(s[6:11].rjust(5) for s in sc))
order = sorted(range(len(data)), key=lambda i: -int(data[i][ix["# Samples"]]))[:topn]
for i in sorted(order):
    r = data[i]
    print(f"{i:5d}", r[ix["# Samples"]].rjust(7), r[ix["Instructions Executed"]].rjust(8), " ".join(r[ix[s]].rjust(5) for s in sc),
          r[ix["Source"]].strip()[:80])
