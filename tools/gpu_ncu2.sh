#!/bin/bash
# ncu --set full capture of one config's sweep kernel: tools/gpu_ncu2.sh <tag> <cfg> [k]
tag=$1; shift
CMD="python tools/one_kernel.py $*"
$CMD > gpurun_out/plain_$tag.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:sweep_ -s 8 -c 2 \
    -o gpurun_out/prof_$tag -f $CMD > gpurun_out/ncu_full_$tag.log 2>&1
echo "$tag rc=$?"
