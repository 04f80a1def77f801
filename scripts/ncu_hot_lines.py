#!/usr/bin/env python
"""Hot source lines of one kernel from `ncu --page source --csv --print-source cuda,sass`."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
agg, hdr, cur = [], None, None
for r in rows:
    if r and r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif r and r[0] == "Line No":
        hdr = r
    elif r and r[0].isdigit() and hdr:
        try:
            inst = int(r[hdr.index("Instructions Executed")])
            thr = int(r[hdr.index("Thread Instructions Executed")])
            samples = int(r[hdr.index("# Samples")]) if r[hdr.index("# Samples")] != "-" else 0
        except (ValueError, IndexError):
            continue
        stalls = {}
        for k, h in enumerate(hdr):
            if h.startswith("stall_") and "Not Issued" not in h and k < len(r):
                try:
                    stalls[h[6:]] = int(r[k])
                except ValueError:
                    pass
        agg.append((cur, int(r[0]), r[1].strip()[:80], inst, thr, samples, stalls))
tot_i = sum(a[3] for a in agg) or 1
tot_s = sum(a[5] for a in agg) or 1
print(f"total warp-inst {tot_i}, samples {tot_s}")
allst = {}
for a in agg:
    for k, v in a[6].items():
        allst[k] = allst.get(k, 0) + v
print("stalls:", ", ".join(f"{k} {v / tot_s * 100:.1f}%" for k, v in sorted(allst.items(), key=lambda kv: -kv[1])[:8]))
for a in sorted(agg, key=lambda a: -a[5])[:top]:
    st = ",".join(f"{k}:{v}" for k, v in sorted(a[6].items(), key=lambda kv: -kv[1])[:2] if v)
    print(f"{a[0][:16]:16s}:{a[1]:4d} inst {a[3] / tot_i * 100:5.1f}% thr {a[4] / max(1, a[3]):4.1f} samp {a[5] / tot_s * 100:5.1f}% [{st}] {a[2]}")
