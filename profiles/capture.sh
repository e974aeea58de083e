#!/bin/bash
# profiles/capture.sh <tag> -- the three measurements the profiles/ directory holds, in the order
# /opt/skills/guides/B200_PROFILING.md asks for: (1) the plain bench (numbers come from THIS run
# only), (2) the ncu launch list of the same command, (3) one `--set full` capture of the step
# kernels (first timestep at full cells + one later timestep), exported as raw CSV.
# Run on the GPU box:  gpurun --timeout 1500 -- 'bash profiles/capture.sh r01_v5'
set -u
tag=${1:-scratch}
out=gpurun_out
mkdir -p $out
python bench.py > $out/bench_$tag.json 2> $out/bench_$tag.err || { echo "plain bench failed"; tail -5 $out/bench_$tag.err; exit 1; }
python profiles/summarize_bench.py $out/bench_$tag.json
# (2) launch list: same command, fewer repetitions (ncu serialises and replays every launch)
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv \
    --log-file $out/launches_$tag.csv python bench.py --steps 1 --warmup 3 > $out/ncu_list_$tag.log 2>&1 \
    || { echo "ncu launch list failed"; tail -5 $out/ncu_list_$tag.log; }
# (3) full capture of one run's step kernels: skip the warm-up runs' launches by kernel-name
# filter + launch-skip (the first 2 timesteps of the first warm-up run are as good as any: the
# state is reset before every run)
ncu --set full --clock-control none --import-source on \
    -k regex:'k_force_integrate|k_move|k_collide_cells|k_collide_big|k_resolve|k_com|k_classify|k_sort_gather|k_sort_scatter' \
    -c 24 -o $out/prof_$tag -f python bench.py --steps 1 --warmup 3 > $out/ncu_full_$tag.log 2>&1 \
    || { echo "ncu full capture failed"; tail -5 $out/ncu_full_$tag.log; }
ncu -i $out/prof_$tag.ncu-rep --page raw --csv > $out/prof_${tag}_raw.csv 2>/dev/null
python profiles/ncu_summary.py $out/prof_${tag}_raw.csv
