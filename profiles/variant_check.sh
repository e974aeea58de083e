#!/bin/bash
# profiles/variant_check.sh <tag> [variants...] -- A/B of force-kernel variants (PARSIM_FORCE_VARIANT)
# against the default: per-timestep force times (tools/force_steps.py, about 4 s each) and, for the
# FIRST variant listed, the GPU parity suite under that variant.  Every leg runs under its own
# timeout: a deadlocked pipeline must not hang the box.
#   gpurun --timeout 300 -- 'bash profiles/variant_check.sh r02_v1 12 13 14 9 8 5'
set -u
tag=${1:-scratch}
shift
variants=${*:-"12 13 14 9 8 5"}
out=gpurun_out
mkdir -p $out
for v in 0 $variants; do
    PARSIM_DEBUG=1 PARSIM_FORCE_VARIANT=$v timeout -s KILL 30 python tools/force_steps.py 2>> $out/force_steps_$tag.err
done | tee $out/force_steps_$tag.txt
first=${variants%% *}
( PARSIM_FORCE_VARIANT=$first timeout -s KILL 120 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_multi.py ) \
    > $out/pytest_gpu_${tag}_fv$first.log 2>&1
echo "pytest under variant $first rc=$?"; tail -n 3 $out/pytest_gpu_${tag}_fv$first.log
grep -h "blocks per SM" $out/force_steps_$tag.err | sort -u
