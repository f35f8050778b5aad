#!/bin/bash
# focused N-GPU comparison of the two halo-exchange modes: tools/gpu_p2p8.sh N [tag]
N=${1:-8}; tag=${2:-r01}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
L=gpurun_out/p2p_${tag}_n$N.log
: > $L
timeout 200 $TR --master-port 29611 tools/multi_check.py >> $L 2>&1; echo "multi_check N=$N rc=$?" | tee -a $L
grep -E "MISMATCH|timed out|Error|error" $L | head -10
for ex in p2p nccl; do
  timeout 200 $TR --master-port 29612 tools/config5.py --mode strong --steps 40 --reps 4 --exchange $ex 2>&1 | grep '^{' | tee -a $L
  timeout 200 $TR --master-port 29614 bench.py --gpus $N --steps 4 --warmup 3 --no-cpu --no-e2e --exchange $ex 2>&1 | grep '^{' | cut -c1-330 | tee -a $L
done
