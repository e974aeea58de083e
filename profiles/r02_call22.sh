#!/bin/bash
# round 2, call 22 (one GPU): upload with late velocities -- the merge of the pieces, the centres of mass and the
# force sums of the first timestep run while vx, vy cross PCIe (A/B with PARSIM_VEL_LATE=0)
set -u
tag=${1:-r02_v22}
out=gpurun_out
mkdir -p $out
( timeout -s KILL 900 python -m pytest tests -m gpu -x -q ) > $out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?"; tail -n 3 $out/pytest_gpu_$tag.log
for v in 1 0 1 0; do
  PARSIM_VEL_LATE=$v timeout -s KILL 400 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras > $out/bench_${tag}_late$v.json 2>> $out/bench_$tag.err; echo "bench late=$v rc=$?"
  python profiles/summarize_bench.py $out/bench_${tag}_late$v.json > $out/tmp.txt 2>&1; grep "^value" $out/tmp.txt | cut -c1-220
done
rm -f $out/tmp.txt
tail -n 3 $out/bench_$tag.err
