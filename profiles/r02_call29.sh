#!/bin/bash
# round 2, call 29 (one GPU): evidence set of the round's final code -- GPU suite, smoke, plain bench line,
# reference arm, ncu launch list of the bench command, ncu --set full of the step kernels (timesteps 1-2 of a
# config-3 run; exported to CSV on the box, the .ncu-rep is deleted: the return path takes 64 MiB)
set -u
tag=${1:-r02_v29}
out=gpurun_out
mkdir -p $out
( timeout -s KILL 900 python -m pytest tests -m gpu -x -q ) > $out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?"; tail -n 3 $out/pytest_gpu_$tag.log
python -c "import __graft_entry__ as g; g.smoke()" > $out/smoke_$tag.log 2>&1; echo "smoke rc=$?"; tail -n 2 $out/smoke_$tag.log
timeout -s KILL 600 python bench.py --steps 10 --warmup 3 > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench rc=$?"
python profiles/summarize_bench.py $out/bench_$tag.json | cut -c1-220
timeout -s KILL 600 python bench.py --impl reference --steps 3 --warmup 1 > $out/bench_${tag}_reference.json 2> $out/bench_${tag}_reference.err; echo "reference arm rc=$?"
cut -c1-300 $out/bench_${tag}_reference.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $out/launches_$tag.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras > $out/ncu_list_$tag.log 2>&1 || { echo "ncu launch list failed"; tail -3 $out/ncu_list_$tag.log; }
python profiles/launch_share.py $out/launches_$tag.csv $out/bench_$tag.json > $out/launch_share_$tag.txt 2>&1; head -n 24 $out/launch_share_$tag.txt
timeout 500 ncu --set full --clock-control none --import-source on \
    -k regex:'k_force_integrate|k_move|k_collide_cells|k_resolve|k_com' -s 46 -c 8 \
    -o $out/prof_${tag}_step -f python tools/force_steps.py > $out/ncu_${tag}_step.log 2>&1 || { echo "ncu step failed"; tail -5 $out/ncu_${tag}_step.log; }
ncu -i $out/prof_${tag}_step.ncu-rep --page raw --csv > $out/ncu_step_${tag}_raw.csv 2>/dev/null
rm -f $out/prof_${tag}_step.ncu-rep
python profiles/ncu_summary.py $out/ncu_step_${tag}_raw.csv > $out/ncu_step_${tag}_summary.txt 2>&1; cat $out/ncu_step_${tag}_summary.txt
du -sh $out
