#!/usr/bin/env python
"""Summarise an `ncu --page raw --csv` dump: one line per profiled launch."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
idx = {h: i for i, h in enumerate(hdr)}
cols = [("gpu__time_duration.sum", "us"), ("launch__registers_per_thread", "regs"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("smsp__thread_inst_executed_per_inst_executed.ratio", "thr/inst"),
        ("smsp__inst_executed.sum", "warp-inst"), ("dram__bytes_read.sum", "dram_rd"),
        ("dram__bytes_write.sum", "dram_wr"), ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("lts__t_bytes.sum", "l2_bytes"), ("l1tex__t_bytes.sum", "l1_bytes")]
print("kernel".ljust(34) + "".join(name.rjust(12) for _, name in cols))
for d in data:
    line = d[idx["Kernel Name"]][:33].ljust(34)
    for key, _ in cols:
        v = d[idx[key]] if key in idx else "-"
        u = units[idx[key]] if key in idx else ""
        try:
            f = float(v.replace(",", ""))
            v = f"{f:.3g}" + ({"Mbyte": "M", "Gbyte": "G", "Kbyte": "K", "byte": "B"}.get(u, ""))
        except ValueError:
            pass
        line += v.rjust(12)
    print(line)
