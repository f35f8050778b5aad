#!/bin/bash
# ncu --set full capture of one config's sweep kernel, summarised ON the box (the reports are
# too large to bring back in bulk): tools/gpu_ncu2.sh <tag> <cfg> [k]   (env: SKIP, COUNT)
tag=$1; shift
mkdir -p gpurun_out
CMD="python tools/one_kernel.py $*"
$CMD > gpurun_out/plain_$tag.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:sweep_ -s ${SKIP:-8} -c ${COUNT:-2} \
    -o /tmp/prof_$tag -f $CMD > gpurun_out/ncu_full_$tag.log 2>&1
rc=$?
if [ $rc -eq 0 ]; then
  { echo "# ncu --set full: tools/one_kernel.py $* (tag $tag)"; echo; python tools/ncu_summary.py /tmp/prof_$tag.ncu-rep;
    echo; echo "## SASS instruction mix (share of executed warp instructions)"; echo; echo '```';
    python tools/ncu_mix.py /tmp/prof_$tag.ncu-rep 20; echo '```'; } > gpurun_out/ncu_$tag.md
  ncu -i /tmp/prof_$tag.ncu-rep --page raw --csv > gpurun_out/ncu_${tag}_raw.csv
fi
echo "$tag rc=$rc"
