#!/bin/bash
# profiles/capture_variant.sh <tag> <variant> -- ncu `--set full` of the force kernel of ONE variant
# (PARSIM_FORCE_VARIANT), first full timestep + two thinner ones, next to the same capture of the
# default; raw CSV + the stall table.  Run on the GPU box AFTER the plain A/B (variant_check.sh):
#   gpurun --timeout 600 -- 'bash profiles/capture_variant.sh r02_v1 5'
set -u
tag=${1:-scratch}
variant=${2:-5}
out=gpurun_out
mkdir -p $out
for v in 0 $variant; do  # PARSIM_DEBUG prints the resident blocks per SM of the persistent variants
    PARSIM_DEBUG=1 PARSIM_FORCE_VARIANT=$v timeout 240 ncu --set full --clock-control none --import-source on \
        -k regex:'k_force_integrate|k_force_tiles' -c 3 -o $out/prof_${tag}_fv$v -f \
        python tools/force_steps.py > $out/ncu_${tag}_fv$v.log 2>&1 || { echo "ncu fv$v failed"; tail -5 $out/ncu_${tag}_fv$v.log; }
    ncu -i $out/prof_${tag}_fv$v.ncu-rep --page raw --csv > $out/prof_${tag}_fv${v}_raw.csv 2>/dev/null
    python profiles/ncu_summary.py $out/prof_${tag}_fv${v}_raw.csv
done
