#!/bin/bash
# profiles/r02_call10.sh <tag> -- GPU tests and bench lines with the upload pieces sorted under the
# copy's shadow (e2e), A/B against PARSIM_PIECE_SORT=0.
set -u
tag=${1:-scratch}
out=gpurun_out
mkdir -p $out
( timeout -s KILL 900 python -m pytest tests -m gpu -x -q ) > $out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?"; tail -n 3 $out/pytest_gpu_$tag.log
timeout -s KILL 600 python bench.py --steps 10 --warmup 3 > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench rc=$?"
python profiles/summarize_bench.py $out/bench_$tag.json | grep -v "^ fp64" | cut -c1-220
PARSIM_PIECE_SORT=0 timeout -s KILL 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-extras > $out/bench_${tag}_nopieces.json 2> $out/bench_${tag}_nopieces.err
echo "== PARSIM_PIECE_SORT=0"; python profiles/summarize_bench.py $out/bench_${tag}_nopieces.json | sed -n 1p | cut -c1-120
timeout -s KILL 300 python bench.py --steps 5 --warmup 3 --workload config4w --no-cpu-baseline > $out/bench_${tag}_config4w.json 2> $out/bench_${tag}_config4w.err
python profiles/summarize_bench.py $out/bench_${tag}_config4w.json | sed -n 1p | cut -c1-120
du -sh $out
