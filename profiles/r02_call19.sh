#!/bin/bash
# round 2, call 19 (one GPU): two-launch double scan on grids of few tiles (A/B with PARSIM_SCAN_FEW=0)
set -u
tag=${1:-r02_v19}
out=gpurun_out
mkdir -p $out
( timeout -s KILL 900 python -m pytest tests -m gpu -x -q ) > $out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?"; tail -n 3 $out/pytest_gpu_$tag.log
for v in 1 0 1 0; do
PARSIM_SCAN_FEW=$v timeout -s KILL 300 python bench.py --steps 20 --warmup 5 --workload config2 --no-cpu-baseline --no-extras > $out/bench_${tag}_config2_few$v.json 2>> $out/bench_${tag}.err; echo "config2 few=$v rc=$?"
python profiles/summarize_bench.py $out/bench_${tag}_config2_few$v.json > $out/tmp.txt 2>&1; grep "^value\|^scan\|^ghost" $out/tmp.txt | cut -c1-150
done
rm -f $out/tmp.txt
