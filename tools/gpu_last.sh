#!/bin/bash
# Short closing pass (few GPU-minutes left): bench line first, then the parity tests that
# compare the CUDA path with the reference-produced golden grids.
mkdir -p gpurun_out
timeout 170 python bench.py > gpurun_out/bench_f.json 2> gpurun_out/bench_f.err; echo "bench rc=$?"
tail -c 600 gpurun_out/bench_f.json
timeout 120 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "reference" > gpurun_out/pytest_f.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/pytest_f.log
