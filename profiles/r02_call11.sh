#!/bin/bash
# round 2, call 11: the allocation-free entry point alone in a fresh process, then the whole GPU suite
tag=${1:-r02_v11}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader > $out/${tag}_gpu.txt
python -m pytest tests/test_zz_no_alloc.py -x -q > $out/${tag}_noalloc_alone.log 2>&1; echo "alone rc=$?" | tee -a $out/${tag}_noalloc_alone.log
python -m pytest tests -m gpu -x -q > $out/${tag}_pytest_gpu.log 2>&1; echo "suite rc=$?" | tee -a $out/${tag}_pytest_gpu.log
tail -n 5 $out/${tag}_noalloc_alone.log $out/${tag}_pytest_gpu.log
