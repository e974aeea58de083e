#!/bin/bash
# round 2, call 26 (one GPU): k_resolve finds the lanes of a cell as a run (shuffle + ballot) instead of match.any;
# A/B against the previous build at config 3 and at 8 per cell
set -u
tag=${1:-r02_v26}
out=gpurun_out
mkdir -p $out
( timeout -s KILL 900 python -m pytest tests -m gpu -x -q ) > $out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?"; tail -n 3 $out/pytest_gpu_$tag.log
for wl in config3 config4w; do
for lib in new prev new prev; do
  if [ $lib = prev ]; then export PARSIM_LIB=$PWD/tools/abtest/lib_prev.so; else unset PARSIM_LIB; fi
  timeout -s KILL 300 python bench.py --steps 5 --warmup 3 --workload $wl --no-cpu-baseline --no-extras > $out/bench_${tag}_${wl}_$lib.json 2>> $out/bench_${tag}.err; echo "$wl $lib rc=$?"
  python profiles/summarize_bench.py $out/bench_${tag}_${wl}_$lib.json > $out/tmp.txt 2>&1; grep "^value\|^collide_resolve" $out/tmp.txt | cut -c1-150
done
done
unset PARSIM_LIB
rm -f $out/tmp.txt
