#!/usr/bin/env python3
"""Collision count and particle 0 of one full-size tuple, for tests/golden/bench_pins.json.

    python tools/pin_run.py <seed> <side> <ncside> <n_part> <n_step> [n_strips]

Runs the tuple through the C ABI on `n_strips` row strips of THIS process (parsim_create: one GPU
per strip when the box has them, else oversubscribed) with device-side init, and prints one JSON
line.  Used where the unmodified serial reference cannot pin a weak-scaling tuple itself: 8e8
particles need more RAM than the build container has, and on some tuples the reference dies of its
own out-of-grid store (serial/parsim.cpp:276-280).  Such pins say so in their `source`."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from particlesimulation_b200 import ParticleSimulation  # noqa: E402

seed, side, nc, n, steps = int(sys.argv[1]), float(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
strips = int(sys.argv[6]) if len(sys.argv) > 6 else 1
t0 = time.time()
sim = ParticleSimulation(seed, side, nc, n, n_gpus=strips, device_init=True)
t1 = time.time()
sim.simulate(steps)
x0, y0, col = sim.result()
print(json.dumps({"args": f"{seed} {side:g} {nc} {n} {steps}", "strips": strips, "transport": os.environ.get("PARSIM_TRANSPORT", "auto"),
                  "x0": f"{x0:.3f}", "y0": f"{y0:.3f}", "collisions": col, "init_s": round(t1 - t0, 2),
                  "run_s": round(time.time() - t1, 3)}), flush=True)
sim.close()
