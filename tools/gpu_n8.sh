#!/bin/bash
# 8-GPU pass (gpurun --gpus 8 -- bash tools/gpu_n8.sh): PCIe fabric probe, bench at N=8 (weak headline,
# config-5 strong leg, decomposed e2e through one call), the C++ CLI on config 5, the device-group tests.
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_n8.txt 2>&1
lscpu | grep -E "NUMA|Socket|Model name|^CPU\(s\)" > gpurun_out/lscpu_n8.txt 2>&1
timeout 300 tools/pcie_probe 512 > gpurun_out/pcie_probe_n8.log 2>&1; echo "probe rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; echo "bench rc=$?"
timeout 600 stencil_b200/host/build/bin/stencil_main -s 1 -i 8 -b 1 --shape 1024,1024,2048 --dtype f64 --init random \
    --gpus 8 -R 2 -m CUDA -c > gpurun_out/cli_config5_n8.log 2>&1; echo "cli rc=$?"
timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/pytest_multi_n8.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/pytest_multi_n8.log; tail -8 gpurun_out/cli_config5_n8.log; cat gpurun_out/pcie_probe_n8.log
