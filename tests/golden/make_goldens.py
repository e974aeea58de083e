#!/usr/bin/env python3
"""Generate parity fixtures from the UNMODIFIED reference serial binary.

TEST INFRASTRUCTURE.  Runs oracle/_ref/parsim (built by oracle/Makefile from
/root/reference/serial/parsim.cpp with the reference's own flags, serial/Makefile:1-3) on the
CLI tuples given on the command line (or the default list) and appends one JSON line per run
to tests/golden/reference_runs.jsonl:

    {"args": [seed, side, ncside, n_part, n_step], "x0": "...", "y0": "...",
     "collisions": N, "seconds": t, "host": "...", "binary": "serial"}

x0/y0 are kept as the printed 3-decimal strings (serial/parsim.cpp:450-453).  The dense
regime (BASELINE.json configs 2-5) is not covered by the reference's own vectors
(serial/run_tests.sh:30-42 are all sparse), so these runs are the pins for it.

usage: make_goldens.py ["seed side ncside n_part n_step" ...]
"""
import json, os, platform, subprocess, sys, time

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "..", "..", "oracle", "_ref", "parsim")
OUT = os.path.join(HERE, "reference_runs.jsonl")

DEFAULT = [
    "1 2 3 10 1",
    "1 10 100 1000000 20",      # BASELINE config 2
    "7 10 100 1000000 10",      # config 3 at 1/100 size (same density / cell size)
    "3 100 1000 8000000 10",    # config 4 at 1/100 size
    "5 1 20 100000 50",         # config 5 at 1/100 size
    "5 1 20 1000000 50",        # config 5 at 1/10 size
    "1 10 100 100000 5",
    "5 1 20 20000 5",
    "-7 10 50 200000 10",       # normal (skewed) distribution, dense
]

def run(args):
    t = time.time()
    p = subprocess.run([REF] + args.split(), capture_output=True, text=True, check=True)
    dt = time.time() - t
    lines = p.stdout.strip().splitlines()
    x0, y0 = lines[-2].split()
    rec = {"args": args.split(), "x0": x0, "y0": y0, "collisions": int(lines[-1]),
           "seconds": round(dt, 2), "host": platform.processor() or platform.machine(),
           "binary": "serial", "stdout_extra_lines": len(lines) - 2}
    with open(OUT, "a") as f:
        f.write(json.dumps(rec) + "\n")
    print(rec, flush=True)

if __name__ == "__main__":
    for a in (sys.argv[1:] or DEFAULT):
        run(a)
