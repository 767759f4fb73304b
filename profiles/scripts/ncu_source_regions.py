"""Stall samples of the first kernel of an `ncu -i X.ncu-rep --page source --csv` export, in windows of N SASS instructions
(where in the instruction stream the time goes, and why).  Usage: python profiles/scripts/ncu_source_regions.py src.csv [window]"""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
win = int(sys.argv[2]) if len(sys.argv) > 2 else 100
hdr, data, n = None, [], 0
for r in rows:
    if r and r[0] == "Kernel Name":
        n += 1
        if n > 1:
            break
        continue
    if n == 1 and hdr is None:
        hdr = r
        continue
    if n == 1 and r and r[0].startswith("0x"):
        data.append(r)
ix = {h: i for i, h in enumerate(hdr)}
sc = [h for h in hdr if h.startswith("stall_") and "Not" not in h]
S = lambda i: int(data[i][ix["# Samples"]])
tot = sum(S(i) for i in range(len(data)))
print("total samples", tot, "instructions", len(data))
print("first  samples  share  max-executed  top stall reasons | top opcodes")
for a in range(0, len(data), win):
    b = min(a + win, len(data))
    t = sum(S(i) for i in range(a, b))
    c = collections.Counter()
    for i in range(a, b):
        for s in sc:
            try:
                c[s[6:]] += int(data[i][ix[s]])
            except ValueError:
                pass
    ops = collections.Counter(re.split(r"[ .]", re.sub(r"^@!?U?P\w+\s+", "", data[i][ix["Source"]].strip()))[0] for i in range(a, b))
    ex = max(int(data[i][ix["Instructions Executed"]]) for i in range(a, b))
    print(f"{a:5d} {t:7d} {100 * t / tot:5.1f}% {ex:9d}  {c.most_common(3)} | {ops.most_common(4)}")
