#!/bin/bash
# round 2, call 15 (2-GPU box): one host round trip less per multi-strip step, the count round trip hidden
# behind the ghost-row exchange, raw H2D ceiling in the bench line; whole GPU suite, bench at 1 and 2 GPUs
set -u
tag=${1:-r02_v15}
out=gpurun_out
mkdir -p $out
( timeout -s KILL 900 python -m pytest tests -m gpu -x -q ) > $out/${tag}_pytest_gpu.log 2>&1; echo "suite rc=$?"; tail -n 3 $out/${tag}_pytest_gpu.log
( timeout -s KILL 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29519 \
    bench.py --gpus 2 --steps 10 --warmup 3 ) > $out/${tag}_bench_n2.json 2> $out/${tag}_bench_n2.err; echo "bench n2 rc=$?"
( timeout -s KILL 400 python bench.py --steps 10 --warmup 3 ) > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench n1 rc=$?"
python - <<PY
import json
for f in ("$out/${tag}_bench_n2.json", "$out/${tag}_bench.json"):
    try:
        d = json.loads(open(f).read().strip().split("\n")[-1])
        print(f, {k: d[k] for k in ("value", "ms_per_step", "n_gpus")}, d["e2e"]["ms_per_step"], d["e2e"]["h2d_GBps_slowest_rank"], d["e2e"]["h2d_raw_GBps_slowest_rank"], d["result"]["parity_ok"])
        print("  extra", {k: (v.get("ms_per_run"), v["result"].get("parity_ok")) for k, v in d.get("extra", {}).items()})
    except Exception as e:
        print(f, "no bench line:", e)
PY
tail -n 3 $out/${tag}_bench_n2.err $out/${tag}_bench.err
