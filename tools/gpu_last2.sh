#!/bin/bash
# Closing verification: the whole GPU suite and smoke() on the final tree.
mkdir -p gpurun_out
timeout 200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_f.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_f.log
tail -3 gpurun_out/pytest_gpu_f.log
timeout 90 python __graft_entry__.py smoke > gpurun_out/smoke_f.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke_f.log
tail -7 gpurun_out/smoke_f.log
