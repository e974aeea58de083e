#!/bin/bash
# profiles/r02_call8.sh <tag> -- GPU tests and plain bench lines after the occupancy fixes
# (collision search back at 32 registers, k_com_v without spills, plain scans on huge grids).
set -u
tag=${1:-scratch}
out=gpurun_out
mkdir -p $out
( timeout -s KILL 900 python -m pytest tests -m gpu -x -q ) > $out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?"; tail -n 3 $out/pytest_gpu_$tag.log
timeout -s KILL 600 python bench.py --steps 10 --warmup 3 > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench rc=$?"
python profiles/summarize_bench.py $out/bench_$tag.json | grep -v "^ fp64" | cut -c1-220
timeout -s KILL 300 python bench.py --steps 5 --warmup 3 --workload config4w --no-cpu-baseline > $out/bench_${tag}_config4w.json 2> $out/bench_${tag}_config4w.err; echo "bench config4w rc=$?"
python profiles/summarize_bench.py $out/bench_${tag}_config4w.json | grep -v "^ fp64" | cut -c1-220
for l in 32 16; do
  PARSIM_COM_LANES=$l timeout -s KILL 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-extras > $out/bench_${tag}_com$l.json 2> $out/bench_${tag}_com$l.err
  echo "== PARSIM_COM_LANES=$l"; python profiles/summarize_bench.py $out/bench_${tag}_com$l.json | grep -E "^value|^com " | cut -c1-100
done
du -sh $out
