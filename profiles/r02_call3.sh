#!/bin/bash
# profiles/r02_call3.sh <tag> -- GPU tests (CUDA-graph replay of small problems, search fused into the
# resolve kernel, early-neighbour force variants), force-variant A/B, plain bench lines.
set -u
tag=${1:-scratch}
out=gpurun_out
mkdir -p $out
( timeout -s KILL 900 python -m pytest tests -m gpu -x -q ) > $out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?"; tail -n 6 $out/pytest_gpu_$tag.log
for v in 0 5 6 7; do
    PARSIM_FORCE_VARIANT=$v timeout -s KILL 60 python tools/force_steps.py 2>> $out/force_steps_$tag.err
done | tee $out/force_steps_$tag.txt
timeout -s KILL 600 python bench.py --steps 5 --warmup 3 > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench rc=$?"
python profiles/summarize_bench.py $out/bench_$tag.json
timeout -s KILL 300 python bench.py --steps 3 --warmup 3 --workload config4w --no-cpu-baseline > $out/bench_${tag}_config4w.json 2> $out/bench_${tag}_config4w.err; echo "bench config4w rc=$?"
python profiles/summarize_bench.py $out/bench_${tag}_config4w.json
PARSIM_FUSE_SEARCH=0 timeout -s KILL 300 python bench.py --steps 3 --warmup 3 --workload config4w --no-cpu-baseline > $out/bench_${tag}_config4w_unfused.json 2> $out/bench_${tag}_config4w_unfused.err
python profiles/summarize_bench.py $out/bench_${tag}_config4w_unfused.json 2>/dev/null | sed -n 1p
for g in 1 0; do
    PARSIM_GRAPHS=$g timeout -s KILL 300 python bench.py --steps 10 --warmup 5 --workload config2 --no-cpu-baseline > $out/bench_${tag}_config2_graphs$g.json 2> $out/bench_${tag}_config2_graphs$g.err
    python profiles/summarize_bench.py $out/bench_${tag}_config2_graphs$g.json 2>/dev/null | sed -n 1p
done
du -sh $out
