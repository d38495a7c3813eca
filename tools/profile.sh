#!/bin/bash
# Round profile recipe (run on the GPU box through gpurun): launch list + full capture of the loop kernels.
# usage: tools/profile.sh <tag>
tag=${1:-r01}
out=gpurun_out
python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $out/bench_pre_ncu_$tag.json 2> $out/bench_pre_ncu_$tag.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/launches_$tag.csv \
    python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $out/ncu_launch_$tag.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'ts_rollout|wgrad' -s 12 -c 8 -o $out/full_$tag -f \
    python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $out/ncu_full_$tag.log 2>&1
ncu -i $out/full_$tag.ncu-rep --page raw --csv > $out/ncu_full_$tag.csv 2>/dev/null
ls -la $out | tail -8
