"""Aggregate `msm_trace g<group> c=<c> <step> <ms> ms` lines (stderr of a B2G16_MSM_TRACE=1 run) by step."""
import sys
from collections import OrderedDict
agg = OrderedDict()
for line in open(sys.argv[1]):
    p = line.split()
    if len(p) == 6 and p[0] == "msm_trace":
        c, t = agg.get(p[3], (0, 0.0))
        agg[p[3]] = (c + 1, t + float(p[4]))
tot = sum(t for _, t in agg.values())
print("total %.3f ms" % tot)
for k, (c, t) in agg.items():
    print("%8.3f ms %5.1f%% x%-3d %s" % (t, 100 * t / tot, c, k))
