#!/bin/bash
# round 2, call 25 (one GPU): FP32 pre-filter in the candidate loop of k_collide_cells; A/B against the previous build
set -u
tag=${1:-r02_v25}
out=gpurun_out
mkdir -p $out
( timeout -s KILL 900 python -m pytest tests -m gpu -x -q ) > $out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?"; tail -n 3 $out/pytest_gpu_$tag.log
for lib in new prev new prev; do
  if [ $lib = prev ]; then export PARSIM_LIB=$PWD/tools/abtest/lib_prev.so; else unset PARSIM_LIB; fi
  timeout -s KILL 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras > $out/bench_${tag}_$lib.json 2>> $out/bench_${tag}.err; echo "config3 $lib rc=$?"
  python profiles/summarize_bench.py $out/bench_${tag}_$lib.json > $out/tmp.txt 2>&1; grep "^value\|^collide_find\|^collide_big" $out/tmp.txt | cut -c1-150
done
unset PARSIM_LIB
timeout -s KILL 300 python bench.py --steps 3 --warmup 2 --workload config5 --no-cpu-baseline --no-extras > $out/bench_${tag}_c5.json 2>> $out/bench_${tag}.err; python profiles/summarize_bench.py $out/bench_${tag}_c5.json 2>&1 | grep "^value" | cut -c1-150
rm -f $out/tmp.txt
