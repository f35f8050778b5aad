#!/bin/bash
# ncu evidence for the bench command: launch list + one full capture of the sweep kernel.
# Usage: tools/gpu_ncu.sh <tag> [extra bench args]
tag=${1:-r01}; shift
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e $*"
$CMD > gpurun_out/plain_$tag.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/launches_$tag.csv $CMD > gpurun_out/ncu_list_$tag.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:sweep_ -s 60 -c 2 \
    -o gpurun_out/prof_$tag -f $CMD > gpurun_out/ncu_full_$tag.log 2>&1
echo "rc=$?"
tail -2 gpurun_out/plain_$tag.log
tail -3 gpurun_out/ncu_full_$tag.log
