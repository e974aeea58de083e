#!/bin/bash
# round 2, call 20 (one GPU): per-particle collision search with the arrival-bag head and the cell index
# loaded one dependent round earlier; A/B against the previous build (tools/abtest/lib_prev.so) at 8 per cell
set -u
tag=${1:-r02_v20}
out=gpurun_out
mkdir -p $out
( timeout -s KILL 900 python -m pytest tests -m gpu -x -q ) > $out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?"; tail -n 3 $out/pytest_gpu_$tag.log
for lib in new prev new prev; do
  if [ $lib = prev ]; then export PARSIM_LIB=$PWD/tools/abtest/lib_prev.so; else unset PARSIM_LIB; fi
  timeout -s KILL 300 python bench.py --steps 5 --warmup 3 --workload config4w --no-cpu-baseline --no-extras > $out/bench_${tag}_c4_$lib.json 2>> $out/bench_${tag}.err; echo "config4w $lib rc=$?"
  python profiles/summarize_bench.py $out/bench_${tag}_c4_$lib.json > $out/tmp.txt 2>&1; grep "^value\|^collide_find\|^collide_resolve" $out/tmp.txt | cut -c1-150
done
unset PARSIM_LIB
rm -f $out/tmp.txt
