#!/usr/bin/env python3
"""Summarise an `ncu --page raw --csv` dump: one line per launch with the metrics that matter here."""
import csv, sys, re
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
def col(name):
    return hdr.index(name) if name in hdr else None
W = [('gpu__time_duration.sum', 'us', 1e-3 if True else 1), ('dram__bytes_read.sum', 'rdMB', 1), ('dram__bytes_write.sum', 'wrMB', 1),
     ('launch__registers_per_thread', 'regs', 1), ('sm__warps_active.avg.pct_of_peak_sustained_active', 'occ%', 1),
     ('smsp__issue_active.avg.pct_of_peak_sustained_active', 'issue%', 1),
     ('sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active', 'fp64%', 1),
     ('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'dram%', 1),
     ('l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed', 'smem%', 1),
     ('smsp__inst_executed.sum', 'Minst', 1e-6)]
print(f"{'kernel':28s}" + ''.join(f"{w[1]:>9s}" for w in W))
for r in rows[2:]:
    name = re.sub(r'\(.*', '', r[col('Kernel Name')])[:28]
    out = f"{name:28s}"
    for m, lab, sc in W:
        i = col(m)
        if i is None: out += f"{'-':>9s}"; continue
        v = float(r[i].replace(',', ''))
        u = units[i]
        if m == 'gpu__time_duration.sum':
            v = v * {'ns': 1e-3, 'us': 1, 'ms': 1e3, 's': 1e6}.get(u, 1)
        elif 'bytes' in m:
            v = v * {'byte': 1e-6, 'Kbyte': 1e-3, 'Mbyte': 1, 'Gbyte': 1e3}.get(u, 1)
        else:
            v *= sc
        out += f"{v:9.1f}"
    print(out)
