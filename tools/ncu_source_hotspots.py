#!/usr/bin/env python3
"""Aggregate `ncu --page source --csv` output: stall samples by code region and the hottest instructions.

    ncu -i gpurun_out/prof_X.ncu-rep --page source --csv > /tmp/x.csv ; python tools/ncu_source_hotspots.py /tmp/x.csv
"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
chunk = int(sys.argv[2]) if len(sys.argv) > 2 else 300
hdr, data = rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
agg = {h: 0 for h in stalls}
reg = {}
tot = 0
for k, r in enumerate(data):
    if not r[ix["# Samples"]].isdigit():
        continue
    s = int(r[ix["# Samples"]])
    tot += s
    e = reg.setdefault(k // chunk, [0, {}])
    e[0] += s
    for h in stalls:
        v = int(r[ix[h]]) if r[ix[h]].isdigit() else 0
        agg[h] += v
        e[1][h] = e[1].get(h, 0) + v
print("kernel:", rows[0][1][:100])
print("total samples", tot, "instructions", len(data))
print("top stalls:", sorted(((v, h) for h, v in agg.items()), reverse=True)[:8])
for region, (s, d) in sorted(reg.items()):
    print(f"  instr {region * chunk:5d}+: {s:7d}", sorted(((v, h[6:]) for h, v in d.items()), reverse=True)[:4])
hot = sorted(((int(r[ix["# Samples"]]), k, r[ix["Source"]][:70]) for k, r in enumerate(data)
              if r[ix["# Samples"]].isdigit()), reverse=True)[:20]
for h in hot:
    print("  hot", h)
