#!/bin/bash
# profiles/final_check.sh <tag> [ngpu] -- one short GPU call, most important leg first (every leg
# writes its own file under gpurun_out/, so a call that runs out of time still brings back what
# finished): multi-strip parity tests (when the box has >= 2 GPUs), the plain 1-GPU bench line,
# the whole GPU test suite, the N-GPU bench line.
set -u
tag=${1:-scratch}
ngpu=${2:-1}
out=gpurun_out
mkdir -p $out
if [ "$ngpu" -gt 1 ]; then
    ( time timeout 200 python -m pytest tests/test_gpu_multi.py -m gpu -x -q ) > $out/pytest_gpu_multi_$tag.log 2>&1; echo "pytest multi rc=$?"; tail -3 $out/pytest_gpu_multi_$tag.log
fi
( time python bench.py ) > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench rc=$?"
python profiles/summarize_bench.py $out/bench_$tag.json 2>&1 | head -3
( time timeout 200 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_multi.py ) > $out/pytest_gpu_$tag.log 2>&1; echo "pytest rc=$?"; tail -3 $out/pytest_gpu_$tag.log
if [ "$ngpu" -gt 1 ]; then
    ( time timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $ngpu --master-addr 127.0.0.1 --master-port 29517 \
        bench.py --gpus $ngpu --steps 3 --warmup 3 ) > $out/bench_${tag}_n$ngpu.json 2> $out/bench_${tag}_n$ngpu.err; echo "bench n$ngpu rc=$?"
    python profiles/summarize_bench.py $out/bench_${tag}_n$ngpu.json 2>&1 | head -3
fi
ls -la $out | tail -12
