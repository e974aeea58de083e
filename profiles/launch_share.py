#!/usr/bin/env python3
"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list, next to the
per-kernel CUDA-event times of the plain bench run (bench JSON `kernels`).  The ncu times are
cold-cache and serialised, so the SHARES are what must agree, not the absolute numbers.
usage: launch_share.py launches.csv [bench.json]"""
import csv, json, re, sys
from collections import defaultdict

GROUP = [("k_force_integrate", "force_integrate"), ("k_collide_cells", "collide_find"), ("k_collide_find", "collide_find"),
         ("k_collide_big", "collide_big_cells"), ("k_collide_huge", "collide_big_cells"), ("k_move", "merge_move"),
         ("k_resolve", "collide_resolve"), ("k_sort_gather", "sort_gather"), ("k_sort_hist", "sort_passes"),
         ("k_sort_scatter", "sort_passes"), ("k_sort_cell_of", "bin_keys"), ("k_com", "com"), ("k_classify", "classify_movers"),
         ("k_scan", "scan"), ("k_ghost", "ghost_rows"), ("k_add_counts", "tombstones")]
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr = rows[0]
kn, mv, mu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot, cnt = defaultdict(float), defaultdict(int)
for r in rows[1:]:
    name = re.sub(r"^void ", "", r[kn])
    name = re.sub(r"[<(].*", "", name)
    ns = float(r[mv].replace(",", "")) * {"ns": 1, "us": 1e3, "ms": 1e6}.get(r[mu], 1)
    tot[name] += ns
    cnt[name] += 1
step = {k: v for k, v in tot.items() if not k.startswith(("k_init", "k_fp64", "k_pair", "k_ingest"))}
s = sum(step.values())
bench = json.load(open(sys.argv[2]))["kernels"] if len(sys.argv) > 2 else None
bs = sum(v["ms_per_step"] for v in bench.values()) if bench else 0
grp = defaultdict(float)
for k, v in step.items():
    g = next((b for a, b in GROUP if k.startswith(a)), k)
    grp[g] += v
print(f"{'kernel':22s} {'launches':>8s} {'ncu ms':>9s} {'share':>7s}")
for k in sorted(step, key=lambda k: -step[k]):
    print(f"{k:22s} {cnt[k]:8d} {step[k] / 1e6:9.3f} {step[k] / s:7.3f}")
print(f"{'total (step kernels)':22s} {sum(cnt[k] for k in step):8d} {s / 1e6:9.3f}")
if bench:
    print(f"\n{'group':22s} {'ncu share':>9s} {'bench share':>11s}   (bench: CUDA events of the plain run, {bs:.2f} ms per run)")
    for k in sorted(grp, key=lambda k: -grp[k]):
        b = bench.get(k, {}).get("ms_per_step", 0.0)
        print(f"{k:22s} {grp[k] / s:9.3f} {b / bs:11.3f}")
