#!/bin/bash
# round 2, call 39 (one GPU): ncu --set full of the step kernels at 8 particles per cell (BASELINE configs[3]'s density,
# 1e8 particles on one GPU) with the round's last code, timesteps 1-2; CSV exported on the box
set -u
tag=${1:-r02_v39}
out=gpurun_out
mkdir -p $out
FORCE_STEPS_WORKLOAD="3 354 3536 100000000" timeout -s KILL 120 python tools/force_steps.py > $out/force_steps_${tag}_config4w.txt 2>&1; cat $out/force_steps_${tag}_config4w.txt
FORCE_STEPS_WORKLOAD="3 354 3536 100000000" timeout 500 ncu --set full --clock-control none --import-source on \
    -k regex:'k_force_integrate|k_move|k_collide_find|k_resolve|k_com|k_classify' -s 57 -c 10 \
    -o $out/prof_${tag}_c4 -f python tools/force_steps.py > $out/ncu_${tag}_c4.log 2>&1 || { echo "ncu failed"; tail -5 $out/ncu_${tag}_c4.log; }
ncu -i $out/prof_${tag}_c4.ncu-rep --page raw --csv > $out/ncu_config4density_${tag}_raw.csv 2>/dev/null
rm -f $out/prof_${tag}_c4.ncu-rep
python profiles/ncu_summary.py $out/ncu_config4density_${tag}_raw.csv > $out/ncu_config4density_${tag}_summary.txt 2>&1; cat $out/ncu_config4density_${tag}_summary.txt
du -sh $out
