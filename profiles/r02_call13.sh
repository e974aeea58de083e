#!/bin/bash
# round 2, call 13: merge-kernel centre of mass staged in shared memory (PARSIM_FUSE_COM=2 vs 1 vs 0);
# L2 fetch granularity hint (PARSIM_L2_FETCH=32/128) on the random 64-byte gather of the step-0 sort
tag=${1:-r02_v13}
out=gpurun_out
mkdir -p $out
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "centres_of_mass or stage_by_stage or deferred" > $out/${tag}_pytest_com.log 2>&1; echo "com tests rc=$?"; tail -n 2 $out/${tag}_pytest_com.log
run() { name=$1; shift; env "$@" python bench.py --steps 10 --warmup 3 --no-extras $EXTRA > $out/${tag}_$name.json 2>> $out/${tag}_bench.err; echo "$name rc=$?"; }
EXTRA=""
run c3_fuse1 PARSIM_FUSE_COM=1
run c3_fuse2 PARSIM_FUSE_COM=2
run c3_l2f32 PARSIM_L2_FETCH=32
run c3_l2f128 PARSIM_L2_FETCH=128
EXTRA="--workload config4w"
run c4_fuse1 PARSIM_FUSE_COM=1
run c4_fuse2 PARSIM_FUSE_COM=2
run c4_l2f32 PARSIM_L2_FETCH=32
for f in c3_fuse1 c3_fuse2 c3_l2f32 c3_l2f128 c4_fuse1 c4_fuse2 c4_l2f32; do echo "== $f"; python profiles/summarize_bench.py $out/${tag}_$f.json 2>&1 | cut -c1-200 | grep -v "^ fp64" | head -n 14; done > $out/${tag}_summary.txt
grep "^value\|^==" $out/${tag}_summary.txt | cut -c1-100
grep "L2 fetch" $out/${tag}_bench.err | sort | uniq -c
