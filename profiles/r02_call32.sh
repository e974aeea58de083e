#!/bin/bash
# round 2, call 32 (2-GPU box): bench.py with the stricter end-to-end check (x0, y0 and the collision count of the
# uploaded run equal the device-resident run's bit for bit) at 1 and 2 GPUs
set -u
tag=${1:-r02_v32}
out=gpurun_out
mkdir -p $out
( timeout -s KILL 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-extras ) > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench n1 rc=$?"
( timeout -s KILL 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 \
    bench.py --gpus 2 --steps 5 --warmup 3 --no-extras ) > $out/bench_${tag}_n2.json 2> $out/bench_${tag}_n2.err; echo "bench n2 rc=$?"
python - <<PY
import json
for f in ("$out/bench_$tag.json", "$out/bench_${tag}_n2.json"):
    try:
        d = json.loads(open(f).read().strip().split("\n")[-1])
        print(f, d["ms_per_step"], d["e2e"]["ms_per_step"], d["result"])
    except Exception as e:
        print(f, "no bench line:", e)
PY
tail -n 4 $out/bench_$tag.err $out/bench_${tag}_n2.err
