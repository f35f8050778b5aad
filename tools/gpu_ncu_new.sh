#!/bin/bash
# ncu --set full of the radius and 1-D kernels (summaries only come back)
for c in rad2 rad3f64 line linef64; do SKIP=4 COUNT=1 bash tools/gpu_ncu2.sh $c $c; done
