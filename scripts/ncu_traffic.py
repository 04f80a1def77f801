#!/usr/bin/env python
"""profiles/traffic.json from an `ncu --page raw --csv` dump: DRAM bytes (read + write) per launch
of each wavefront kernel, averaged over the captured launches.  bench.py copies the dominant
kernel's figure into roofline.traffic."""
import csv
import json
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
idx = {h: i for i, h in enumerate(hdr)}
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
out = {}
for d in data:
    name = d[idx["Kernel Name"]]
    key = next((k for k in ("raygen", "extend", "shade", "connect") if k + "_kernel" in name), None)
    if key is None:
        continue
    total = 0.0
    for metric in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        total += float(d[idx[metric]].replace(",", "")) * scale[units[idx[metric]]]
    e = out.setdefault(key, {"dram_bytes_per_launch": 0.0, "launches": 0, "us_under_ncu": 0.0})
    e["dram_bytes_per_launch"] += total
    e["us_under_ncu"] += float(d[idx["gpu__time_duration.sum"]].replace(",", ""))
    e["launches"] += 1
for e in out.values():
    e["dram_bytes_per_launch"] = round(e["dram_bytes_per_launch"] / e["launches"])
    e["us_under_ncu"] = round(e["us_under_ncu"] / e["launches"], 1)
out["_source"] = sys.argv[2] if len(sys.argv) > 2 else sys.argv[1]
json.dump(out, sys.stdout, indent=1)
print()
