#!/bin/bash
# round 2, call 31 (4-GPU box): bench.py under torchrun at 4 GPUs with the collective whole-strip upload
set -u
tag=${1:-r02_v28}
out=gpurun_out
mkdir -p $out
( timeout -s KILL 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus 4 --steps 5 --warmup 3 ) > $out/bench_${tag}_n4.json 2> $out/bench_${tag}_n4.err
echo "bench n8 rc=$?"
python - <<PY
import json
try:
    d = json.loads(open("$out/bench_${tag}_n4.json").read().strip().split("\n")[-1])
    print({k: d[k] for k in ("value", "ms_per_step", "n_gpus")}, d["e2e"], d["result"]["parity_ok"])
    print("config4", d.get("extra", {}).get("config4", {}).get("ms_per_run"), d.get("extra", {}).get("config4", {}).get("result", {}).get("parity_ok"))
except Exception as e:
    print("no bench line:", e)
PY
tail -n 5 $out/bench_${tag}_n4.err
