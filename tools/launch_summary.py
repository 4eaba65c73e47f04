"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total ms,
share.  Usage: python tools/launch_summary.py launches.csv [last_n_launches]"""
import csv
import sys
from collections import OrderedDict

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
scale = {"ns": 1e-6, "nsecond": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0}
L = [(r[ki].split("(")[0].replace("void ", "").replace("b2::", ""), float(r[vi].replace(",", "")) * scale[r[ui]]) for r in rows[1:]]
if len(sys.argv) > 2:
    L = L[-int(sys.argv[2]):]
agg = OrderedDict()
for k, ms in L:
    c, t = agg.get(k, (0, 0.0))
    agg[k] = (c + 1, t + ms)
tot = sum(t for _, t in agg.values())
print("launches %d, total %.3f ms" % (len(L), tot))
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%8.3f ms %5.1f%% x%-3d %s" % (t, 100 * t / tot, c, k[:100]))
