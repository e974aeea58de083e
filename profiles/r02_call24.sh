#!/bin/bash
# round 2, call 24 (2-GPU box): last check of the final code over NCCL -- whole GPU suite, bench at 2 GPUs and 1 GPU
set -u
tag=${1:-r02_v24}
out=gpurun_out
mkdir -p $out
( timeout -s KILL 900 python -m pytest tests -m gpu -x -q ) > $out/pytest_gpu_$tag.log 2>&1; echo "suite rc=$?"; tail -n 3 $out/pytest_gpu_$tag.log
( timeout -s KILL 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 \
    bench.py --gpus 2 --steps 10 --warmup 3 ) > $out/bench_${tag}_n2.json 2> $out/bench_${tag}_n2.err; echo "bench n2 rc=$?"
( timeout -s KILL 400 python bench.py --steps 10 --warmup 3 ) > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench n1 rc=$?"
python - <<PY
import json
for f in ("$out/bench_${tag}_n2.json", "$out/bench_$tag.json"):
    try:
        d = json.loads(open(f).read().strip().split("\n")[-1])
        print(f, {k: d[k] for k in ("value", "ms_per_step", "n_gpus")}, d["e2e"]["ms_per_step"], d["e2e"]["h2d_GBps_slowest_rank"], d["e2e"]["h2d_raw_GBps_slowest_rank"], d["result"]["parity_ok"])
        print("  extra", {k: (v.get("ms_per_run"), v["result"].get("parity_ok")) for k, v in d.get("extra", {}).items()})
    except Exception as e:
        print(f, "no bench line:", e)
PY
tail -n 3 $out/bench_${tag}_n2.err $out/bench_$tag.err
