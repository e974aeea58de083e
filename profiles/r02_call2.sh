#!/bin/bash
# profiles/r02_call2.sh <tag> -- GPU tests with the sync-free step loop, plain bench lines, then ncu
# --set full of (a) the default force kernel with per-instruction stall samples (step 1 and step 3 of
# a config-3 run) and (b) every step kernel at config 4's density (8 per cell).  The .ncu-rep files
# are exported to CSV ON THE BOX and deleted: gpurun_out/ only comes back below 64 MiB.
set -u
tag=${1:-scratch}
out=gpurun_out
mkdir -p $out
( timeout -s KILL 900 python -m pytest tests -m gpu -x -q ) > $out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?"; tail -n 4 $out/pytest_gpu_$tag.log
timeout -s KILL 600 python bench.py --steps 5 --warmup 3 > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench rc=$?"
python profiles/summarize_bench.py $out/bench_$tag.json
PARSIM_SYNC_STEPS=1 timeout -s KILL 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-extras > $out/bench_${tag}_syncsteps.json 2> $out/bench_${tag}_syncsteps.err
python profiles/summarize_bench.py $out/bench_${tag}_syncsteps.json | head -1
timeout -s KILL 300 python bench.py --steps 3 --warmup 3 --workload config4w --no-cpu-baseline > $out/bench_${tag}_config4w.json 2> $out/bench_${tag}_config4w.err; echo "bench config4w rc=$?"
python profiles/summarize_bench.py $out/bench_${tag}_config4w.json | head -1
# (a) force kernel, default instantiation: launches 11 and 13 of tools/force_steps.py = timesteps 1 and 3
timeout 400 ncu --set full --clock-control none --import-source on -k regex:'k_force_integrate' -s 10 -c 3 \
    -o $out/prof_${tag}_force -f python tools/force_steps.py > $out/ncu_${tag}_force.log 2>&1 || { echo "ncu force failed"; tail -5 $out/ncu_${tag}_force.log; }
ncu -i $out/prof_${tag}_force.ncu-rep --page raw --csv > $out/prof_${tag}_force_raw.csv 2>/dev/null
ncu -i $out/prof_${tag}_force.ncu-rep --page source --csv > $out/prof_${tag}_force_source.csv 2>/dev/null
rm -f $out/prof_${tag}_force.ncu-rep
python profiles/ncu_summary.py $out/prof_${tag}_force_raw.csv
# (b) config 4's density: the step kernels of timesteps 1-2
timeout 500 ncu --set full --clock-control none \
    -k regex:'k_force_integrate|k_move|k_collide_find|k_collide_cells|k_resolve|k_com|k_classify' -s 14 -c 14 \
    -o $out/prof_${tag}_c4 -f python bench.py --steps 1 --warmup 1 --workload config4w --no-cpu-baseline > $out/ncu_${tag}_c4.log 2>&1 || { echo "ncu c4 failed"; tail -5 $out/ncu_${tag}_c4.log; }
ncu -i $out/prof_${tag}_c4.ncu-rep --page raw --csv > $out/prof_${tag}_c4_raw.csv 2>/dev/null
rm -f $out/prof_${tag}_c4.ncu-rep
python profiles/ncu_summary.py $out/prof_${tag}_c4_raw.csv
du -sh $out
