#!/bin/bash
# First GPU pass: parity tests, smoke, timing probe, bench; then (only if all of that
# exited 0) the ncu launch list and one full capture of the sweep kernel.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
tail -5 gpurun_out/smoke.log
timeout 900 python tools/gpu_probe.py c2 c1 c4 c3 > gpurun_out/probe.log 2>&1; echo "probe rc=$?" | tee -a gpurun_out/probe.log
cat gpurun_out/probe.log
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench rc=$?" | tee -a gpurun_out/bench.log
tail -3 gpurun_out/bench.log
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2>&1; echo "bench_ref rc=$?"
tail -2 gpurun_out/bench_ref.log
