#!/bin/bash
# profiles/final_check.sh <tag> -- one short GPU call: smoke, the GPU parity tests, the plain bench line,
# the L2-fetch-granularity A/B (PARSIM_L2_FETCH, step-0 gather), the reference arm, then the ncu launch
# list of the bench command.  Every leg writes its own file under gpurun_out/, most important first,
# so a call that runs out of time still brings back what finished.
set -u
tag=${1:-scratch}
out=gpurun_out
mkdir -p $out
( time python __graft_entry__.py smoke ) > $out/smoke_$tag.log 2>&1; echo "smoke rc=$?"
( time timeout 240 python -m pytest tests -m gpu -x -q ) > $out/pytest_gpu_$tag.log 2>&1; echo "pytest rc=$?"; tail -3 $out/pytest_gpu_$tag.log
( time python bench.py ) > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench rc=$?"
python profiles/summarize_bench.py $out/bench_$tag.json 2>&1 | tail -25
for gran in 32 64 128; do
    PARSIM_L2_FETCH=$gran python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $out/bench_${tag}_l2fetch$gran.json 2> $out/bench_${tag}_l2fetch$gran.err
    python - $out/bench_${tag}_l2fetch$gran.json $gran <<'P'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    k = d["kernels"]
    print("l2fetch", sys.argv[2], "ms/run %.2f" % d["ms_per_step"], " ".join("%s=%.2f" % (n, k[n]["ms_per_step"]) for n in ("sort_gather", "sort_passes", "collide_find", "collide_resolve", "merge_move", "classify_movers", "com", "force_integrate") if n in k))
except Exception as e:
    print("l2fetch", sys.argv[2], "failed:", e)
P
done
( time python bench.py --impl reference --steps 2 --warmup 1 ) > $out/bench_${tag}_reference.json 2> $out/bench_${tag}_reference.err; echo "reference rc=$?"
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv \
    --log-file $out/launches_$tag.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $out/ncu_list_$tag.log 2>&1 \
    || { echo "ncu launch list failed"; tail -5 $out/ncu_list_$tag.log; }
ls -la $out | tail -20
