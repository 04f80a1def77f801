#!/usr/bin/env python
"""profiles/traffic.json from an `ncu --page raw --csv` dump: DRAM bytes (read + write) per launch
of each wavefront kernel, averaged over the captured launches.  bench.py copies the dominant
kernel's figure into roofline.traffic."""
import csv
import json
import sys

# usage: ncu_traffic.py raw.csv "source note" [raw2.csv "source note 2" ...]; kernels on the wide BVH get the keys
# extend_bvh / connect_bvh, with the L2 -> L1 read bytes (lts__t_sectors_srcunit_tex_op_read x 32 B) next to the DRAM bytes
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
out = {}
sources = []
for path, note in zip(sys.argv[1::2], sys.argv[2::2]):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    sources.append(note)
    for d in data:
        name = d[idx["Kernel Name"]]
        key = next((k for k in ("raygen", "extend", "shade", "connect") if k + "_kernel" in name), None)
        if key is None:
            continue
        if "WideBvh" in name:
            key += "_bvh"
        total = 0.0
        for metric in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            total += float(d[idx[metric]].replace(",", "")) * scale[units[idx[metric]]]
        e = out.setdefault(key, {"dram_bytes_per_launch": 0.0, "l2_read_bytes_per_launch": 0.0, "launches": 0, "us_under_ncu": 0.0})
        e["dram_bytes_per_launch"] += total
        e["l2_read_bytes_per_launch"] += 32.0 * float(d[idx["lts__t_sectors_srcunit_tex_op_read.sum"]].replace(",", ""))
        e["us_under_ncu"] += float(d[idx["gpu__time_duration.sum"]].replace(",", ""))
        e["launches"] += 1
for e in out.values():
    e["dram_bytes_per_launch"] = round(e["dram_bytes_per_launch"] / e["launches"])
    e["l2_read_bytes_per_launch"] = round(e["l2_read_bytes_per_launch"] / e["launches"])
    e["us_under_ncu"] = round(e["us_under_ncu"] / e["launches"], 1)
    e["dram_gbs_under_ncu"] = round(e["dram_bytes_per_launch"] / e["us_under_ncu"] / 1e3, 1)
    e["l2_read_gbs_under_ncu"] = round(e["l2_read_bytes_per_launch"] / e["us_under_ncu"] / 1e3, 1)
out["_source"] = "; ".join(sources)
json.dump(out, sys.stdout, indent=1)
print()
