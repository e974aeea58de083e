#!/usr/bin/env python3
"""Per-timestep device time of the force + integration launch (config 3), for the variant chosen with
PARSIM_FORCE_VARIANT: tells whether a variant wins in the full first step (100 per cell) or in the
thinner later steps (about 30 per cell).  usage: [PARSIM_FORCE_VARIANT=v] python tools/force_steps.py"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from particlesimulation_b200 import ParticleSimulation  # noqa: E402

# FORCE_STEPS_WORKLOAD="<seed> <side> <ncside> <n_part>" picks another workload (default: config 3)
w = os.environ.get("FORCE_STEPS_WORKLOAD", "7 100 1000 100000000").split()
sim = ParticleSimulation(int(w[0]), float(w[1]), int(w[2]), int(w[3]), device_init=True, options={"keep_initial": 1})
sim.simulate(10)  # warm-up
sim.reset()
sim.set_option("stats", 1)
out = []
for _ in range(10):
    sim.stats_reset()
    sim.simulate(1)
    k = sim.stats()["kernels"]
    out.append(round(k["force_integrate"]["ms"], 3))
print(json.dumps({"variant": os.environ.get("PARSIM_FORCE_VARIANT", "default"), "workload": " ".join(w), "force_ms_per_step": out,
                  "total": round(sum(out), 2), "collisions": sim.result()[2]}))
