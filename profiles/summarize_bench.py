#!/usr/bin/env python3
"""Print the per-kernel table of a bench.py JSON line (profiles/bench_*.json)."""
import json, sys
d = json.load(open(sys.argv[1]))
print(f"value {d['value']:.4g}  ms/run {d['ms_per_step']:.2f}  e2e {d['e2e']['value']:.4g} ({d['e2e']['ms_per_step']:.1f} ms)"
      f"  launches {d['gpu_launches']}  clocks {d['clocks'].get('sm_mhz')}  parity {d['result']}")
for k, v in d.get('fp64', {}).items():
    print(' fp64', k, v)
for k, v in sorted(d['kernels'].items(), key=lambda kv: -kv[1]['ms_per_step']):
    print(f"{k:18s} {v['ms_per_step']:8.3f} ms  L={v['launches_per_step']:5.1f} {v['algorithmic_GB_per_step']:7.2f} GB"
          f" {v['achieved_GBps']:8.1f} GB/s {v['hbm_frac']:.3f}")
for k, v in d.get('extra', {}).items():
    print(f" extra {k}: {v['ms_per_run']:.3f} ms/run  {v['value']:.4g}  {v['result']}")
r = d.get('roofline', {})
print(' roofline', {k: r.get(k) for k in ('kernel', 'bound', 'achieved', 'peak', 'unit', 'frac', 'traffic')})
