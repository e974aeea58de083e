#!/bin/bash
# round 2, call 14 (8-GPU box): strip tests over NCCL / peer copies with the final code, then bench.py under
# torchrun at 8 GPUs (weak scaling at config 3's density + extra.config4 = BASELINE configs[3] at full size)
#   gpurun --gpus 8 --timeout 420 -- 'bash profiles/r02_call14.sh r02_v14'
set -u
tag=${1:-r02_v14}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -n 8
( timeout -s KILL 150 python -m pytest tests/test_gpu_multi.py -m gpu -x -q ) > $out/${tag}_pytest_multi8.log 2>&1
echo "pytest multi rc=$?"; tail -n 3 $out/${tag}_pytest_multi8.log
( timeout -s KILL 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus 8 --steps 5 --warmup 3 ) > $out/${tag}_bench_n8.json 2> $out/${tag}_bench_n8.err
echo "bench n8 rc=$?"
python - <<PY
import json
try:
    d = json.loads(open("$out/${tag}_bench_n8.json").read().strip().split("\n")[-1])
    print({k: d[k] for k in ("value", "ms_per_step", "n_gpus")}, d["e2e"], d["result"])
    print("config4", d.get("extra", {}).get("config4"))
except Exception as e:
    print("no bench line:", e)
PY
tail -n 5 $out/${tag}_bench_n8.err
