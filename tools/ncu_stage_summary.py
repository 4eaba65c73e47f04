"""Per-kernel aggregate of an `ncu --page raw --csv` dump: launches, total ms, DRAM bytes, time-weighted SM / DRAM
throughput, registers.  Usage: python tools/ncu_stage_summary.py raw.csv out.json "description" """
import csv
import json
import sys
from collections import OrderedDict

rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
ix = {h: i for i, h in enumerate(hdr)}
scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1.0, "us": 1e-3, "ns": 1e-6, "s": 1e3, "msecond": 1.0, "usecond": 1e-3, "nsecond": 1e-6, "second": 1e3}


def val(r, k):
    return float(r[ix[k]].replace(",", "")) * scale.get(units[ix[k]], 1.0)


agg = OrderedDict()
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    name = r[ix["Kernel Name"]].split("(")[0].replace("void ", "")
    a = agg.setdefault(name, {"launches": 0, "ms": 0.0, "dram_read_bytes": 0.0, "dram_write_bytes": 0.0, "sm_pct_x_ms": 0.0, "dram_pct_x_ms": 0.0, "issue_pct_x_ms": 0.0,
                              "registers": int(float(r[ix["launch__registers_per_thread"]]))})
    ms = val(r, "gpu__time_duration.sum")
    a["launches"] += 1; a["ms"] += ms
    a["dram_read_bytes"] += val(r, "dram__bytes_read.sum"); a["dram_write_bytes"] += val(r, "dram__bytes_write.sum")
    a["sm_pct_x_ms"] += ms * float(r[ix["sm__throughput.avg.pct_of_peak_sustained_elapsed"]])
    a["dram_pct_x_ms"] += ms * float(r[ix["gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"]])
    a["issue_pct_x_ms"] += ms * float(r[ix["smsp__issue_active.avg.pct_of_peak_sustained_active"]])
out = {"what": sys.argv[3] if len(sys.argv) > 3 else "", "kernels": [], "total_ms": 0.0, "total_dram_bytes": 0.0}
for name, a in sorted(agg.items(), key=lambda kv: -kv[1]["ms"]):
    k = {"kernel": name, "launches": a["launches"], "ms": round(a["ms"], 4), "dram_bytes": a["dram_read_bytes"] + a["dram_write_bytes"],
         "sm_throughput_pct": round(a["sm_pct_x_ms"] / a["ms"], 1), "dram_throughput_pct": round(a["dram_pct_x_ms"] / a["ms"], 1),
         "issue_active_pct": round(a["issue_pct_x_ms"] / a["ms"], 1), "registers": a["registers"]}
    out["kernels"].append(k)
    out["total_ms"] += a["ms"]; out["total_dram_bytes"] += k["dram_bytes"]
json.dump(out, open(sys.argv[2], "w"), indent=1)
print(json.dumps({"total_ms": out["total_ms"], "total_dram_GB": out["total_dram_bytes"] / 1e9, "top": out["kernels"][:4]}))
