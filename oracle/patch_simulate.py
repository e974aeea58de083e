#!/usr/bin/env python3
"""TEST INFRASTRUCTURE ONLY.  Applies INTEGRATION.md section 2's source-level patch to the
reference's serial program and writes the patched translation unit to STDOUT (it is piped straight
into the compiler by oracle/Makefile: no reference source is stored in the repo).

    patch_simulate.py <reference serial/parsim.cpp> <simulate_b200.inc>

The patch: `#include "parsim_b200.h"` after the reference's own includes, and the body of
ParticleSimulation::simulate (serial/parsim.cpp:422-448) replaced by simulate_b200.inc.
"""
import re
import sys

src = open(sys.argv[1]).read()
body = open(sys.argv[2]).read()
m = re.search(r"^[ \t]*void simulate\(long n_time_steps\)\s*\{", src, flags=re.M)
if not m:
    sys.exit("simulate(long n_time_steps) not found in the reference")
depth, i = 1, m.end()
while depth:
    depth += {"{": 1, "}": -1}.get(src[i], 0)
    i += 1
patched = src[:m.start()] + body + src[i:]
last_inc = list(re.finditer(r"^#include .*$", patched, flags=re.M))[-1]
patched = patched[:last_inc.end()] + '\n#include "parsim_b200.h"' + patched[last_inc.end():]
sys.stdout.write(patched)
