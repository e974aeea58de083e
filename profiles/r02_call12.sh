#!/bin/bash
# round 2, call 12: centre of mass fused into the merge (A/B with PARSIM_FUSE_COM=0), merge of the upload
# pieces with preloaded bounds (A/B with PARSIM_MERGE_PRELOAD=0); whole GPU suite first
tag=${1:-r02_v12}
out=gpurun_out
mkdir -p $out
python -m pytest tests -m gpu -x -q > $out/${tag}_pytest_gpu.log 2>&1; echo "suite rc=$?" | tee -a $out/${tag}_pytest_gpu.log
tail -n 3 $out/${tag}_pytest_gpu.log
python bench.py --steps 10 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench rc=$?"
PARSIM_FUSE_COM=0 python bench.py --steps 10 --warmup 3 --no-extras > $out/${tag}_bench_nofuse.json 2>> $out/${tag}_bench.err; echo "bench nofuse rc=$?"
PARSIM_MERGE_PRELOAD=0 python bench.py --steps 10 --warmup 3 --no-extras > $out/${tag}_bench_nopreload.json 2>> $out/${tag}_bench.err; echo "bench nopreload rc=$?"
python bench.py --steps 10 --warmup 3 --no-extras --workload config4w > $out/${tag}_config4w.json 2>> $out/${tag}_bench.err; echo "config4w rc=$?"
PARSIM_FUSE_COM=0 python bench.py --steps 10 --warmup 3 --no-extras --workload config4w > $out/${tag}_config4w_nofuse.json 2>> $out/${tag}_bench.err; echo "config4w nofuse rc=$?"
for f in bench bench_nofuse bench_nopreload config4w config4w_nofuse; do echo "== $f"; python profiles/summarize_bench.py $out/${tag}_$f.json 2>&1 | cut -c1-200 | head -n 16; done > $out/${tag}_summary.txt; head -n 3 $out/${tag}_summary.txt
tail -n 5 $out/${tag}_bench.err
