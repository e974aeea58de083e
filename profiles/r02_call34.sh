#!/bin/bash
# round 2, call 34 (one GPU): wall time of the drop-in CLI at BASELINE configs[2], device init (default) vs host stream
tag=${1:-r02_v34}
out=gpurun_out
mkdir -p $out
TIMEFORMAT="wall %R s"
for m in default 0 default; do
  if [ $m = 0 ]; then export PARSIM_DEVICE_INIT=0; else unset PARSIM_DEVICE_INIT; fi
  echo "PARSIM_DEVICE_INIT=$m:"; ( time particlesimulation_b200/parsim 7 100 1000 100000000 10 ) 2>&1 | tr '\n' ' '; echo
done 2>&1 | tee $out/cli_wall_$tag.txt
