#!/bin/bash
# round 2, call 33 (one GPU): the CLI generates non-negative seeds on the device by default; whole GPU suite,
# and the drop-in's wall time at BASELINE configs[2] with both init modes
set -u
tag=${1:-r02_v33}
out=gpurun_out
mkdir -p $out
( timeout -s KILL 900 python -m pytest tests -m gpu -x -q ) > $out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?"; tail -n 3 $out/pytest_gpu_$tag.log
for m in default 0; do
  if [ $m = 0 ]; then export PARSIM_DEVICE_INIT=0; else unset PARSIM_DEVICE_INIT; fi
  /usr/bin/time -f "init=$m wall %e s" particlesimulation_b200/parsim 7 100 1000 100000000 10 2>&1 | tr '\n' ' '; echo
done 2>&1 | tee $out/cli_wall_$tag.txt
