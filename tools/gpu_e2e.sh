#!/bin/bash
# e2e pass: pipelined-path parity tests, then tools/e2e_probe.py over parallelogram counts
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "pipelined or dropin or host" > gpurun_out/pytest_e2e.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_e2e.log
for c in ${CHUNKS:-auto 5 7 9 11 13 17}; do
  if [ $c = auto ]; then timeout 120 python tools/e2e_probe.py 2>&1 | tail -1
  else SB200_PIPELINE_CHUNKS=$c timeout 120 python tools/e2e_probe.py 2>&1 | tail -1; fi
done | tee gpurun_out/e2e_probe.log
