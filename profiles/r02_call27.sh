#!/bin/bash
# round 2, call 27 (2-GPU box): collective whole-strip upload with late velocities (parsim_upload_strip) --
# GPU suite (incl. bench under torchrun), bench at 2 GPUs with A/B PARSIM_VEL_LATE=0, bench at 1 GPU
set -u
tag=${1:-r02_v27}
out=gpurun_out
mkdir -p $out
( timeout -s KILL 900 python -m pytest tests -m gpu -x -q ) > $out/pytest_gpu_$tag.log 2>&1; echo "suite rc=$?"; tail -n 5 $out/pytest_gpu_$tag.log
for v in 1 0; do
( PARSIM_VEL_LATE=$v timeout -s KILL 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2952$v \
    bench.py --gpus 2 --steps 10 --warmup 3 ) > $out/bench_${tag}_n2_late$v.json 2> $out/bench_${tag}_n2_late$v.err; echo "bench n2 late=$v rc=$?"
done
python - <<PY
import json
for f in ("$out/bench_${tag}_n2_late1.json", "$out/bench_${tag}_n2_late0.json"):
    try:
        d = json.loads(open(f).read().strip().split("\n")[-1])
        print(f, {k: d[k] for k in ("value", "ms_per_step", "n_gpus")}, d["e2e"]["ms_per_step"], d["e2e"]["h2d_GBps_slowest_rank"], d["e2e"]["h2d_raw_GBps_slowest_rank"], d["result"]["parity_ok"])
    except Exception as e:
        print(f, "no bench line:", e)
PY
tail -n 5 $out/bench_${tag}_n2_late1.err
