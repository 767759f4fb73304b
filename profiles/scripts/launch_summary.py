"""Per-kernel and per-class totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list (plain or .gz), as the markdown
tables under profiles/.  One step = the launches from the second-to-last `fbank_kernel` launch of the capture up to the last one
(a whole warm step of `bench.py --graph-rows 0`: the frontend opens every step).
Usage: python profiles/scripts/launch_summary.py launches.csv[.gz]"""
import collections
import csv
import gzip
import re
import sys

path = sys.argv[1]
f = gzip.open(path, "rt") if path.endswith(".gz") else open(path)
rows = [r for r in csv.reader(l for l in f if l.startswith('"'))]
hdr, rows = rows[0], rows[1:]
ix = {h: i for i, h in enumerate(hdr)}
rows = [r for r in rows if r[ix["Metric Name"]] == "gpu__time_duration.sum"]
fb = [i for i, r in enumerate(rows) if "fbank_kernel" in r[ix["Kernel Name"]]]
assert len(fb) >= 2, "the capture holds fewer than two steps"
rows = rows[fb[-2]:fb[-1]]
n = len(rows)
unit = rows[0][ix["Metric Unit"]]
scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3}[unit.replace("second", "s").replace("n", "n")[:2] if unit not in ("ns", "us", "ms") else unit]
agg = collections.OrderedDict()
for r in rows:
    name = re.sub(r"^void ", "", r[ix["Kernel Name"]])
    name = re.sub(r"\(.*$", "", name)
    k = (name, r[ix["Grid Size"]])
    a = agg.setdefault(k, [0, 0.0])
    a[0] += 1
    a[1] += float(r[ix["Metric Value"]].replace(",", "")) * scale
tot = sum(a[1] for a in agg.values())
print(f"one step = {n} launches, {tot / 1e3:.2f} ms serialised under ncu (cold caches, unthrottled clock)\n")
print("| kernel | grid | launches | mean us | total ms | share |\n|---|---|---|---|---|---|")
for (name, grid), (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"| `{name}` | {grid} | {c} | {t / c:.1f} | {t / 1e3:.3f} | {100 * t / tot:.1f} % |")
cls = collections.Counter()
for (name, _), (c, t) in agg.items():
    key = ("gemm" if "gemm" in name else "attention+fsmn" if "attention" in name or "fsmn" in name else "frontend" if "fbank" in name
           else "layernorm" if "layernorm" in name else "other")
    cls[key] += t
print("\nclasses: " + ", ".join(f"{k} {v / 1e3:.2f} ms ({100 * v / tot:.1f} %)" for k, v in cls.most_common()))
