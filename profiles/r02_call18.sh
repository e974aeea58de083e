#!/bin/bash
# round 2, call 18 (one GPU): one-block 1024-thread scans on small grids, A/B against the three-launch scans
set -u
tag=${1:-r02_v18}
out=gpurun_out
mkdir -p $out
( timeout -s KILL 900 python -m pytest tests -m gpu -x -q ) > $out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?"; tail -n 3 $out/pytest_gpu_$tag.log
for v in 16384 0; do
PARSIM_SCAN_SMALL=$v timeout -s KILL 300 python bench.py --steps 20 --warmup 5 --workload config2 --no-cpu-baseline --no-extras > $out/bench_${tag}_config2_small$v.json 2>> $out/bench_${tag}.err; echo "config2 small=$v rc=$?"
python profiles/summarize_bench.py $out/bench_${tag}_config2_small$v.json > $out/tmp.txt 2>&1; grep "^value\|^scan\|^ghost" $out/tmp.txt | cut -c1-150
done
rm -f $out/tmp.txt
