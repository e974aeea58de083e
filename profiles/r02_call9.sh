#!/bin/bash
# profiles/r02_call9.sh <tag> -- 2-GPU box: the whole GPU suite (strip tests over NCCL and over peer
# copies, error propagation across strips), bench.py at 1 and 2 GPUs.
set -u
tag=${1:-scratch}
out=gpurun_out
mkdir -p $out
( timeout -s KILL 1200 python -m pytest tests -m gpu -x -q ) > $out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?"; tail -n 4 $out/pytest_gpu_$tag.log
timeout -s KILL 600 python bench.py --steps 10 --warmup 3 > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench rc=$?"
python profiles/summarize_bench.py $out/bench_$tag.json | grep -v "^ fp64" | cut -c1-220
timeout -s KILL 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 \
    bench.py --gpus 2 --steps 5 --warmup 3 > $out/bench_${tag}_n2.json 2> $out/bench_${tag}_n2.err
echo "bench n=2 rc=$?"; python profiles/summarize_bench.py $out/bench_${tag}_n2.json | grep -v "^ fp64" | cut -c1-220 | head -12
timeout -s KILL 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 \
    bench.py --impl reference --gpus 2 --steps 1 --warmup 0 2> $out/bench_${tag}_n2_reference.err | cut -c1-300
du -sh $out
