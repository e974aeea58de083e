#!/bin/bash
# round 2, call 30 (one GPU): k_move<.., COM> with the new list's end bound loaded up front; A/B against the previous build
set -u
tag=${1:-r02_v30}
out=gpurun_out
mkdir -p $out
for wl in config3 config4w; do
for lib in new prev new prev; do
  if [ $lib = prev ]; then export PARSIM_LIB=$PWD/tools/abtest/lib_prev.so; else unset PARSIM_LIB; fi
  timeout -s KILL 300 python bench.py --steps 5 --warmup 3 --workload $wl --no-cpu-baseline --no-extras > $out/bench_${tag}_${wl}_$lib.json 2>> $out/bench_${tag}.err; echo "$wl $lib rc=$?"
  python profiles/summarize_bench.py $out/bench_${tag}_${wl}_$lib.json > $out/tmp.txt 2>&1; grep "^value\|^merge_move" $out/tmp.txt | cut -c1-150
done
done
unset PARSIM_LIB
rm -f $out/tmp.txt
