"""profiles/r2_hbm_kernels.md from the ncu csv of tools/prof_hbm.py (tools/r3j.sh): per-kernel average duration, DRAM bytes per
launch and the share of the measured copy bandwidth (MEASURED_PEAKS.json).

    python tools/hbm_summary.py gpurun_out/r3j_hbm.csv [note]
"""
import collections
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
h = rows[0]
ki, mi, vi, ui, ii = (h.index(k) for k in ("Kernel Name", "Metric Name", "Metric Value", "Metric Unit", "ID"))
per = collections.OrderedDict()
for r in rows[1:]:
    d = per.setdefault((r[ii], r[ki]), {})
    v = float(r[vi].replace(",", ""))
    if r[mi].startswith("gpu__time"):
        d["ns"] = v * {"ns": 1, "us": 1e3, "ms": 1e6}.get(r[ui], 1)
    else:
        d[r[mi]] = v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(r[ui], 1)
agg = collections.OrderedDict()
for (_, k), d in per.items():
    name = k.split("(")[0].replace("void ", "").replace("drm::", "")[:48]
    a = agg.setdefault(name, [0, 0.0, 0.0])
    a[0] += 1
    a[1] += d.get("ns", 0)
    a[2] += d.get("dram__bytes_read.sum", 0) + d.get("dram__bytes_write.sum", 0)
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
out = ["# r2: HBM-bound kernels of one whole update (walker, 50 x 50 observed, 2 500 x 15 imagined; `tools/prof_hbm.py` under",
       "# `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none`, 3 updates;",
       f"# per-launch times are cold-cache and serialised).  Peak = measured copy bandwidth {peak:.0f} GB/s (MEASURED_PEAKS.json).",
       "# Kernels far below the peak move a few MB per launch: they are launch- / latency-sized, not bandwidth-sized."]
if len(sys.argv) > 2:
    out.append("# " + sys.argv[2])
out += ["", "| kernel | launches | avg us | DRAM MB / launch | achieved GB/s | of peak |", "|---|---|---|---|---|---|"]
for n, (c, ns, b) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    gbs = b / ns if ns else 0
    out.append(f"| {n} | {c} | {ns / c / 1e3:.1f} | {b / c / 1e6:.2f} | {gbs:.0f} | {gbs / peak * 100:.0f} % |")
open(os.path.join(ROOT, "profiles", "r2_hbm_kernels.md"), "w").write("\n".join(out) + "\n")
print("\n".join(out))
