#!/bin/bash
# round 2, call 37 (one GPU): whole GPU suite with the download / dump-right-after-upload checks; the no-alloc test alone
set -u
tag=${1:-r02_v37}
out=gpurun_out
mkdir -p $out
( timeout -s KILL 900 python -m pytest tests -m gpu -x -q ) > $out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?"; tail -n 12 $out/pytest_gpu_$tag.log
( timeout -s KILL 300 python -m pytest tests/test_zz_no_alloc.py -x -q ) > $out/pytest_noalloc_alone_$tag.log 2>&1
echo "noalloc alone rc=$?"; tail -n 2 $out/pytest_noalloc_alone_$tag.log
