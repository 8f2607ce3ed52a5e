"""Turns an `ncu --metrics gpu__time_duration.sum --csv` launch list and an `ncu --set full` report into the tracked
artefacts under profiles/: a per-kernel table (markdown) and profiles/ncu_traffic.json (DRAM bytes per launch of the
rollout kernels, read by bench.py for roofline.traffic).

    python tools/profile_summary.py gpurun_out/r1b_bf16_launches.csv gpurun_out/r1b_bf16_rollout.ncu-rep r1b
"""
import collections
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
launch_csv, rep, tag = sys.argv[1], sys.argv[2], sys.argv[3]

rows = [r for r in csv.reader(open(launch_csv)) if len(r) > 5]
hdr = rows[0]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
agg = collections.OrderedDict()
for r in rows[1:]:
    try:
        v = float(r[vi].replace(",", ""))
    except ValueError:
        continue
    a = agg.setdefault(r[ki][:70], [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
lines = ["| kernel | launches | total us | share |", "|---|---|---|---|"]
for n, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    lines.append(f"| {n} | {a[0]} | {a[1] / 1e3:.1f} | {a[1] / tot * 100:.1f}% |")

raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rr = list(csv.reader(raw.splitlines()))
h = rr[0]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size",
        "launch__registers_per_thread", "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed.avg.per_cycle_active",
        "lts__t_sector_hit_rate.pct"]
seen = {}
full = ["", "| kernel | " + " | ".join(w.split(".")[0] for w in want) + " |", "|---|" + "---|" * len(want)]
traffic = {}
for r in rr[2:]:
    name = r[h.index("Kernel Name")].split("(")[0]
    if name in seen:
        continue
    seen[name] = 1
    vals = [r[h.index(w)] + " " + rr[1][h.index(w)] for w in want]
    full.append(f"| {name} | " + " | ".join(vals) + " |")

    def by(k):
        v, u = float(r[h.index(k)].replace(",", "")), rr[1][h.index(k)]
        return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]

    if "_tc_" in name:  # the tcgen05 imagination rollouts (the fp32 observation rollouts are listed in the table only)
        key = "rollout_fwd" if "fwd" in name else "rollout_bwd"
        traffic[f"walker:2500:bf16:{key}"] = by("dram__bytes_read.sum") + by("dram__bytes_write.sum")
with open(os.path.join(ROOT, "profiles", f"{tag}_kernels.md"), "w") as f:
    f.write(f"# {tag}: launch list ({os.path.basename(launch_csv)}) and ncu --set full of the rollout kernels "
            f"(tools/prof_r2.py: one whole RSSM update, 50 x 50 observed + 2 500 x 15 imagined states)\n\n" + "\n".join(lines) + "\n" + "\n".join(full) + "\n")
with open(os.path.join(ROOT, "profiles", "ncu_traffic.json"), "w") as f:
    json.dump(traffic, f, indent=1)
print("\n".join(lines[:12]))
print(traffic)
