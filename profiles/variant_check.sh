#!/bin/bash
# profiles/variant_check.sh <tag> -- A/B of the persistent force kernel (PARSIM_FORCE_VARIANT=5/6)
# against the default, every leg under its own timeout (a deadlocked pipeline must not hang the box).
set -u
tag=${1:-scratch}
out=gpurun_out
mkdir -p $out
show() { python - "$1" "$2" <<'P'
import json, sys
try:
    d = json.load(open(sys.argv[1])); k = d["kernels"]
    print(sys.argv[2], "ms/run %.2f" % d["ms_per_step"], "force=%.2f" % k["force_integrate"]["ms_per_step"], "parity", d["result"]["parity_ok"], d["result"]["collisions"])
except Exception as e:
    print(sys.argv[2], "failed:", e)
P
}
PARSIM_FORCE_VARIANT=5 timeout -s KILL 50 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $out/bench_${tag}_fv5.json 2> $out/bench_${tag}_fv5.err; echo "fv5 bench rc=$?"
show $out/bench_${tag}_fv5.json fv5
( PARSIM_FORCE_VARIANT=5 timeout -s KILL 80 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_multi.py ) > $out/pytest_gpu_${tag}_fv5.log 2>&1; echo "fv5 pytest rc=$?"; tail -3 $out/pytest_gpu_${tag}_fv5.log
PARSIM_FORCE_VARIANT=6 timeout -s KILL 50 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $out/bench_${tag}_fv6.json 2> $out/bench_${tag}_fv6.err; echo "fv6 bench rc=$?"
show $out/bench_${tag}_fv6.json fv6
tail -2 $out/bench_${tag}_fv5.err $out/bench_${tag}_fv6.err
