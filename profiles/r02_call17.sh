#!/bin/bash
# round 2, call 17 (one GPU): fewer launches per timestep for launch-bound problems (step reset folded into
# k_ghost_wrap, one-block scans on grids of <= 65536 cells); GPU suite, configs[1] and the default line
set -u
tag=${1:-r02_v17}
out=gpurun_out
mkdir -p $out
( timeout -s KILL 900 python -m pytest tests -m gpu -x -q ) > $out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?"; tail -n 3 $out/pytest_gpu_$tag.log
timeout -s KILL 300 python bench.py --steps 20 --warmup 5 --workload config2 --no-cpu-baseline --no-extras > $out/bench_${tag}_config2.json 2> $out/bench_${tag}.err; echo "config2 rc=$?"
python profiles/summarize_bench.py $out/bench_${tag}_config2.json | cut -c1-200 | head -n 16
PARSIM_GRAPHS=0 timeout -s KILL 300 python bench.py --steps 20 --warmup 5 --workload config2 --no-cpu-baseline --no-extras > $out/bench_${tag}_config2_nographs.json 2>> $out/bench_${tag}.err
python profiles/summarize_bench.py $out/bench_${tag}_config2_nographs.json | head -n 1 | cut -c1-120
timeout -s KILL 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $out/bench_$tag.json 2>> $out/bench_$tag.err; echo "bench rc=$?"
python profiles/summarize_bench.py $out/bench_$tag.json | grep "^value\|extra" | cut -c1-200
