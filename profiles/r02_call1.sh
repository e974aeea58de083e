#!/bin/bash
# profiles/r02_call1.sh -- first GPU call of round 2: strip tests on one GPU (device-copy transport),
# force-variant A/B (two targets per thread), plain bench lines, then ncu --set full of the force
# kernel (default, two-target variant 13, persistent TMA variant 5) with source-level stalls.
#   gpurun --timeout 1500 -- 'bash profiles/r02_call1.sh r02_v1'
set -u
tag=${1:-scratch}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv,noheader
timeout -s KILL 120 tools/micro/n3peak > $out/n3peak_$tag.txt 2>&1; echo "n3peak rc=$?"; cat $out/n3peak_$tag.txt
( timeout -s KILL 900 python -m pytest tests -m gpu -x -q ) > $out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?"; tail -n 5 $out/pytest_gpu_$tag.log
for v in 0 12 13 14 5 15; do
    PARSIM_FORCE_VARIANT=$v timeout -s KILL 60 python tools/force_steps.py 2>> $out/force_steps_$tag.err
done | tee $out/force_steps_$tag.txt
( PARSIM_FORCE_VARIANT=15 timeout -s KILL 900 python -m pytest tests -m gpu -x -q ) > $out/pytest_gpu_${tag}_fv15.log 2>&1
echo "pytest under variant 15 (Newton-3 force kernel) rc=$?"; tail -n 5 $out/pytest_gpu_${tag}_fv15.log
timeout -s KILL 600 python bench.py --steps 5 --warmup 3 > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench rc=$?"
python profiles/summarize_bench.py $out/bench_$tag.json
timeout -s KILL 300 python bench.py --steps 3 --warmup 3 --workload config4w --no-cpu-baseline > $out/bench_${tag}_config4w.json 2> $out/bench_${tag}_config4w.err; echo "bench config4w rc=$?"
python profiles/summarize_bench.py $out/bench_${tag}_config4w.json
timeout -s KILL 300 python bench.py --steps 5 --warmup 3 --workload config2 --no-cpu-baseline > $out/bench_${tag}_config2.json 2> $out/bench_${tag}_config2.err; echo "bench config2 rc=$?"
python profiles/summarize_bench.py $out/bench_${tag}_config2.json
for v in 0 13 5 15; do
    PARSIM_FORCE_VARIANT=$v timeout 400 ncu --set full --clock-control none --import-source on \
        -k regex:'k_force_integrate|k_force_tiles|k_force_n3' -s 10 -c 3 -o $out/prof_${tag}_fv$v -f \
        python tools/force_steps.py > $out/ncu_${tag}_fv$v.log 2>&1 || { echo "ncu fv$v failed"; tail -5 $out/ncu_${tag}_fv$v.log; }
    ls -la $out/prof_${tag}_fv$v.ncu-rep
done
