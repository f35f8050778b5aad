#!/bin/bash
# Final scaling pass on an 8-GPU box (fused peer-store exchange): tools/gpu_scale.sh [tag]
tag=${1:-r01d}
mkdir -p gpurun_out
L=gpurun_out/scale_$tag.log
: > $L
for n in 8 4 2; do
  TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $n"
  timeout 200 $TR --master-port $((29800+n)) bench.py --gpus $n --steps 5 --warmup 3 --no-cpu 2>&1 | grep '^{' | tee -a $L | cut -c1-160
  timeout 200 $TR --master-port $((29810+n)) tools/config5.py --mode strong --steps 40 --reps 3 2>&1 | grep '^{' | tee -a $L
done
timeout 200 python bench.py --gpus 1 --steps 5 --warmup 3 --no-cpu 2>&1 | grep '^{' | tee -a $L | cut -c1-160
timeout 200 python tools/config5.py --mode strong --steps 40 --reps 3 2>&1 | grep '^{' | tee -a $L
