#!/bin/bash
# round 2, call 35 (2-GPU box): collective strip upload -- worker tests (result == one GPU, twice on the same contexts;
# a failing rank takes the other one with it instead of leaving it waiting), whole GPU suite, bench at 2 GPUs
set -u
tag=${1:-r02_v35}
out=gpurun_out
mkdir -p $out
( timeout -s KILL 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -k "collective_strip_upload" ) > $out/pytest_upload_$tag.log 2>&1; echo "upload tests rc=$?"; tail -n 15 $out/pytest_upload_$tag.log
( timeout -s KILL 900 python -m pytest tests -m gpu -x -q ) > $out/pytest_gpu_$tag.log 2>&1; echo "suite rc=$?"; tail -n 3 $out/pytest_gpu_$tag.log
( timeout -s KILL 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29543 \
    bench.py --gpus 2 --steps 5 --warmup 3 --no-extras ) > $out/bench_${tag}_n2.json 2> $out/bench_${tag}_n2.err; echo "bench n2 rc=$?"
python - <<PY
import json
try:
    d = json.loads(open("$out/bench_${tag}_n2.json").read().strip().split("\n")[-1])
    print(d["ms_per_step"], d["e2e"]["ms_per_step"], d["result"]["parity_ok"])
except Exception as e:
    print("no bench line:", e)
PY
