#!/bin/bash
# Build the z-wrapped probe library HERE (nvcc cross-compiles), then:
#   gpurun -- 'python tools/sm_bound_probe.py product; python tools/sm_bound_probe.py probe'
set -e
cd "$(dirname "$0")/../stencil_b200/csrc"
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -cudart static \
     -DSB200_PROBE_ZWRAP=8 -c -o /tmp/capi_probe.o capi.cu
nvcc -gencode arch=compute_100a,code=sm_100a -shared -Xcompiler -fPIC -cudart static \
     -o ../libstencil_b200_probe.so /tmp/capi_probe.o resident2d.o warp2d.o tma3db.o
