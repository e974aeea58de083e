#!/bin/bash
# round 2, call 21 (one GPU): force kernel with the cell's sources read through L1 (broadcast loads) instead of
# a staged shared-memory range -- variants 12 (neighbours still staged) and 13 (no barrier at all) against the default
set -u
tag=${1:-r02_v21}
out=gpurun_out
mkdir -p $out
( timeout -s KILL 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "variant" ) > $out/pytest_variants_$tag.log 2>&1
echo "pytest rc=$?"; tail -n 2 $out/pytest_variants_$tag.log
for v in 0 12 13 2; do
  PARSIM_FORCE_VARIANT=$v timeout -s KILL 120 python tools/force_steps.py 2>> $out/force_steps_$tag.err
done | tee $out/force_steps_${tag}.txt
