#!/bin/bash
# profiles/r02_call6.sh <tag> -- GPU tests and the plain bench lines of the final round-2 code
# (particle count by value where the host is in the loop, deferred + graph-replayed steps only for
# launch-bound problems), the ncu launch list of the bench command and the reference arm.
set -u
tag=${1:-scratch}
out=gpurun_out
mkdir -p $out
( timeout -s KILL 900 python -m pytest tests -m gpu -x -q ) > $out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?"; tail -n 4 $out/pytest_gpu_$tag.log
python -c "import __graft_entry__ as g; g.smoke()" > $out/smoke_$tag.log 2>&1; echo "smoke rc=$?"; tail -n 2 $out/smoke_$tag.log
timeout -s KILL 600 python bench.py --steps 10 --warmup 3 > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench rc=$?"
python profiles/summarize_bench.py $out/bench_$tag.json
timeout -s KILL 300 python bench.py --steps 3 --warmup 3 --workload config4w --no-cpu-baseline > $out/bench_${tag}_config4w.json 2> $out/bench_${tag}_config4w.err; echo "bench config4w rc=$?"
python profiles/summarize_bench.py $out/bench_${tag}_config4w.json
timeout -s KILL 300 python bench.py --steps 10 --warmup 5 --workload config2 --no-cpu-baseline > $out/bench_${tag}_config2.json 2> $out/bench_${tag}_config2.err
python profiles/summarize_bench.py $out/bench_${tag}_config2.json 2>/dev/null | sed -n 1p
timeout -s KILL 600 python bench.py --impl reference --steps 3 --warmup 1 > $out/bench_${tag}_reference.json 2> $out/bench_${tag}_reference.err; echo "reference arm rc=$?"
cut -c1-400 $out/bench_${tag}_reference.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $out/launches_$tag.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras > $out/ncu_list_$tag.log 2>&1 || { echo "ncu launch list failed"; tail -3 $out/ncu_list_$tag.log; }
python profiles/launch_share.py $out/launches_$tag.csv $out/bench_$tag.json 2>/dev/null | head -30
du -sh $out
