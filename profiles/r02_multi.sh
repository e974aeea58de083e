#!/bin/bash
# profiles/r02_multi.sh <tag> -- on a 4-GPU box: strip tests over NCCL and over peer copies, the
# full-size tuples of the weak-scaling series on 1 / 2 / 4 strips (pins for bench_pins.json, and
# N strips == 1 GPU at full size), then bench.py under torchrun at 2 and 4 GPUs.
#   gpurun --gpus 4 --timeout 1500 -- 'bash profiles/r02_multi.sh r02_v3'
set -u
tag=${1:-scratch}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=index,name --format=csv,noheader
( timeout -s KILL 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q ) > $out/pytest_multi_$tag.log 2>&1
echo "pytest multi rc=$?"; tail -n 4 $out/pytest_multi_$tag.log
for v in 1 5 6; do  # sparse-cell force variants at config 4's density (one GPU)
    FORCE_STEPS_WORKLOAD="3 354 3536 100000000" PARSIM_FORCE_VARIANT=$v timeout -s KILL 60 python tools/force_steps.py 2>> $out/force_steps_$tag.err
done | tee $out/force_steps_${tag}_config4w.txt
nvidia-smi topo -m > $out/topo_$tag.txt 2>&1; (numactl -H || lscpu | grep -i numa) >> $out/topo_$tag.txt 2>&1
pins=$out/pins_$tag.jsonl
: > $pins
run() { timeout -s KILL 300 python tools/pin_run.py "$@" >> $pins 2>> $out/pins_$tag.err; tail -n 1 $pins; }
run 7 141.4 1414 200000000 10 1      # pinned by the serial reference: 58430961
run 7 141.4 1414 200000000 10 2
run 3 500 5000 200000000 10 1        # the serial reference dies on this tuple
run 3 500 5000 200000000 10 2
run 7 200 2000 400000000 10 1
run 7 200 2000 400000000 10 4
PARSIM_TRANSPORT=local run 7 200 2000 400000000 10 4
run 3 707 7071 400000000 10 1
run 3 707 7071 400000000 10 4
run 7 282.8 2828 800000000 10 4      # 8e8 particles: more than one GPU holds
run 3 1000 10000 800000000 10 4      # BASELINE configs[3] itself
for n in 2 4; do
    timeout -s KILL 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n \
        bench.py --gpus $n --steps 5 --warmup 3 > $out/bench_${tag}_n$n.json 2> $out/bench_${tag}_n$n.err
    echo "bench n=$n rc=$?"; python profiles/summarize_bench.py $out/bench_${tag}_n$n.json | grep -v "^ fp64" | head -8
done
du -sh $out
