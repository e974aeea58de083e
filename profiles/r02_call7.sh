#!/bin/bash
# profiles/r02_call7.sh <tag> -- GPU tests (centre-of-mass lanes follow the occupancy), then an A/B
# of config 4's density between library builds of this round (PARSIM_LIB): which change moved it.
set -u
tag=${1:-scratch}
out=gpurun_out
mkdir -p $out
( timeout -s KILL 900 python -m pytest tests -m gpu -x -q ) > $out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?"; tail -n 4 $out/pytest_gpu_$tag.log
for lib in tools/abtest/libparsim_91c9e25.so tools/abtest/libparsim_19e5369.so particlesimulation_b200/libparsim_b200.so; do
    for rep in 1 2; do
        PARSIM_LIB=$PWD/$lib timeout -s KILL 300 python bench.py --steps 5 --warmup 3 --workload config4w --no-cpu-baseline > $out/ab_$tag.json 2> $out/ab_$tag.err
        echo "== $lib rep $rep"; python profiles/summarize_bench.py $out/ab_$tag.json | grep -E "^value|collide_find|scan |force_integrate|merge_move" | cut -c1-120
    done
done
timeout -s KILL 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench rc=$?"
python profiles/summarize_bench.py $out/bench_$tag.json | grep -v "^ fp64" | cut -c1-200
du -sh $out
