"""Summarise an `ncu --page source --csv` export: stall totals and the hottest SASS instructions."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
idx = {h: i for i, h in enumerate(hdr)}
data = []
for r in rows[2:]:
    if len(r) == len(hdr) and r[idx["# Samples"]].isdigit():
        data.append(r)
    elif data and r and r[0] == "Kernel Name":
        break  # next launch's section
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[idx["# Samples"]]) for r in data)
print("kernel", rows[0][1], "total samples", tot, "instructions", len(data))
agg = {s: sum(int(r[idx[s]]) for r in data) for s in stalls}
print({k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v > 0})
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
for i in sorted(range(len(data)), key=lambda i: -int(data[i][idx["# Samples"]]))[:n]:
    r = data[i]
    st = {s: int(r[idx[s]]) for s in stalls if int(r[idx[s]]) > 0}
    st = dict(sorted(st.items(), key=lambda kv: -kv[1])[:2])
    print(str(i).rjust(5), r[idx["# Samples"]].rjust(6), r[idx["Source"]].strip()[:64].ljust(64), st)
