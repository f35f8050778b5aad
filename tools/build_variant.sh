#!/bin/bash
# tools/build_variant.sh <tag> <extra nvcc flags...> : a second library stencil_b200/libstencil_b200_<tag>.so
# built from the same sources with extra -D flags (kernel experiments; never the product build).
set -e
tag=$1; shift
cd "$(dirname "$0")/../stencil_b200/csrc"
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -cudart static "$@" -c -o /tmp/capi_$tag.o capi.cu
nvcc -gencode arch=compute_100a,code=sm_100a -shared -Xcompiler -fPIC -cudart static -o ../libstencil_b200_$tag.so /tmp/capi_$tag.o resident2d.o warp2d.o tma3db.o
echo built $tag
