#!/usr/bin/env python3
"""Force-kernel FP64 rate on a given config: python tools/force_rate.py seed side nc n steps"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from particlesimulation_b200 import ParticleSimulation
seed, side, nc, n, steps = int(sys.argv[1]), float(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
g = ParticleSimulation(seed, side, nc, n, device_init=seed >= 0, options={"stats": 1, "keep_initial": 1})
g.simulate(steps)  # warm-up
g.reset(); g.stats_reset()
g.simulate(steps)
s = g.stats()
peak = g.fp64_peak_tflops() / 2 * 1e12  # lane-FMAs per second
k = s["kernels"]["force_integrate"]
ops = 13.0 * s["force_pairs"] + 130.0 * s["alive_particle_steps"]
print(f"{sys.argv[1:]}: force {k['ms']:.3f} ms, pairs {s['force_pairs']:.3e}, alive-steps {s['alive_particle_steps']:.3e}, "
      f"fp64 lane-ops/s {ops / (k['ms'] * 1e-3):.3e} = {ops / (k['ms'] * 1e-3) / peak:.3f} of FMA-chain peak")
pp = g.pair_peak()
print(f"   pair-mix peak {pp:.3e} pairs/s; force kernel {s['force_pairs'] / (k['ms'] * 1e-3):.3e} pairs/s (+ epilogue) = {s['force_pairs'] / (k['ms'] * 1e-3) / pp:.3f}")
for name, v in s["kernels"].items():
    if v["launches"]:
        print(f"   {name:18s} {v['ms']:9.3f} ms")
