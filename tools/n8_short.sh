#!/bin/bash
# Short 8-GPU pass (bounded GPU-minutes): bench.py under torchrun, config 4 in y-slabs (fused vs NCCL), group tests.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
N=${1:-8}
timeout 100 $TR --nproc-per-node $N --master-port 29581 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_n${N}x.json 2> gpurun_out/bench_n${N}x.err; echo "bench rc=$?"
for ex in p2p nccl; do timeout 40 $TR --nproc-per-node $N --master-port 29613 tools/config4.py --exchange $ex 2>&1 | grep '^{'; done | tee gpurun_out/config4_n${N}x.log
timeout 60 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -2 | tee gpurun_out/pytest_multi_n${N}x.log
