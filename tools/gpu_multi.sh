#!/bin/bash
# Multi-GPU pass on an N-GPU box: tools/gpu_multi.sh N [tag]
N=${1:-2}; tag=${2:-r01}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
L=gpurun_out/multi_${tag}_n$N.log
: > $L
timeout 300 $TR --nproc-per-node $N --master-port 29511 tools/multi_check.py >> $L 2>&1; echo "multi_check N=$N rc=$?" | tee -a $L
n=$N
while [ $n -ge 1 ]; do
  if [ $n -gt 1 ]; then RUN="$TR --nproc-per-node $n --master-port $((29520+n))"; else RUN="python"; fi
  timeout 300 $RUN bench.py --gpus $n --steps 3 --warmup 3 --no-cpu --no-e2e 2>&1 | grep '^{' | tee -a $L
  timeout 300 $RUN tools/config5.py --mode strong --steps 20 2>&1 | grep '^{' | tee -a $L
  timeout 300 $RUN tools/config5.py --mode weak --steps 20 2>&1 | grep '^{' | tee -a $L
  n=$((n/2))
done
