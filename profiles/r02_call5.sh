#!/bin/bash
# profiles/r02_call5.sh <tag> -- GPU tests (dual scan, one reset launch per step), sparse-variant A/B
# on config 3, plain bench lines, then ncu --set full with stall samples of the collision search and
# the merge kernel of a config-3 run (timesteps 1-2).
set -u
tag=${1:-scratch}
out=gpurun_out
mkdir -p $out
( timeout -s KILL 900 python -m pytest tests -m gpu -x -q ) > $out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?"; tail -n 4 $out/pytest_gpu_$tag.log
for v in 1 5; do
    PARSIM_SPARSE_VARIANT=$v timeout -s KILL 60 python tools/force_steps.py 2>> $out/force_steps_$tag.err
done | tee $out/force_steps_${tag}_sparse_variant.txt
timeout -s KILL 600 python bench.py --steps 5 --warmup 3 > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench rc=$?"
python profiles/summarize_bench.py $out/bench_$tag.json
timeout -s KILL 300 python bench.py --steps 3 --warmup 3 --workload config4w --no-cpu-baseline > $out/bench_${tag}_config4w.json 2> $out/bench_${tag}_config4w.err; echo "bench config4w rc=$?"
python profiles/summarize_bench.py $out/bench_${tag}_config4w.json
timeout -s KILL 300 python bench.py --steps 10 --warmup 5 --workload config2 --no-cpu-baseline > $out/bench_${tag}_config2.json 2> $out/bench_${tag}_config2.err
python profiles/summarize_bench.py $out/bench_${tag}_config2.json 2>/dev/null | sed -n 1p
timeout 400 ncu --set full --clock-control none --import-source on -k regex:'k_collide_cells|k_move|k_resolve|k_com' -s 40 -c 8 \
    -o $out/prof_${tag}_step -f python tools/force_steps.py > $out/ncu_${tag}_step.log 2>&1 || { echo "ncu failed"; tail -5 $out/ncu_${tag}_step.log; }
ncu -i $out/prof_${tag}_step.ncu-rep --page raw --csv > $out/prof_${tag}_step_raw.csv 2>/dev/null
ncu -i $out/prof_${tag}_step.ncu-rep --page source --csv > $out/prof_${tag}_step_source.csv 2>/dev/null
rm -f $out/prof_${tag}_step.ncu-rep
python profiles/ncu_summary.py $out/prof_${tag}_step_raw.csv
du -sh $out
