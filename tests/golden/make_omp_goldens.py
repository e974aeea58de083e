#!/usr/bin/env python3
"""Fixtures for the optional `semantics = omp` mode from the UNMODIFIED
reference OpenMP binary.

TEST INFRASTRUCTURE.  Runs oracle/_ref/parsim-omp (built by oracle/Makefile from
/root/reference/openMP/parsim-omp.cpp with the reference's own flags) with OMP_NUM_THREADS=1 and
appends one JSON line per run to tests/golden/omp_runs.jsonl.  One thread on purpose: with more
threads the order in which leavers are appended to their new cells depends on the schedule, and
the greedy collision count with it (`5 1 20 20000 5` gave 5307, 5328 and 5313 on three runs with
4-8 threads against 5331 with one), so only the one-thread program is a function of its inputs.

usage: make_omp_goldens.py ["seed side ncside n_part n_step" ...]
"""
import json, os, platform, subprocess, sys, time

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "..", "..", "oracle", "_ref", "parsim-omp")
OUT = os.path.join(HERE, "omp_runs.jsonl")

DEFAULT = ["1 2 3 10 1", "1 10 100 100000 5", "5 1 20 20000 5", "-7 10 50 200000 10", "-10 3 3 100 10",
           "1 10 100 1000000 20", "7 10 100 1000000 10", "3 100 1000 8000000 10", "5 1 20 100000 50"]

def run(args):
    t = time.time()
    p = subprocess.run([REF] + args.split(), capture_output=True, text=True, check=True,
                       env=dict(os.environ, OMP_NUM_THREADS="1"))
    lines = p.stdout.strip().splitlines()
    x0, y0 = lines[-2].split()
    rec = {"args": args.split(), "x0": x0, "y0": y0, "collisions": int(lines[-1]),
           "seconds": round(time.time() - t, 2), "host": platform.processor() or platform.machine(),
           "binary": "openmp", "threads": 1}
    with open(OUT, "a") as f:
        f.write(json.dumps(rec) + "\n")
    print(rec, flush=True)

if __name__ == "__main__":
    for a in (sys.argv[1:] or DEFAULT):
        run(a)
