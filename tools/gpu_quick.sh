#!/bin/bash
# quick kernel iteration pass: tools/gpu_quick.sh <tag> <pytest -k expr> <probe cfgs...>
tag=$1; kexpr=$2; shift 2
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "$kexpr" > gpurun_out/pytest_$tag.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_$tag.log
timeout 600 python tools/gpu_probe.py $* > gpurun_out/probe_$tag.log 2>&1; cat gpurun_out/probe_$tag.log
