#!/bin/bash
# Validation of the schedule-driven pipelined host path (build_pipeline + executor).
mkdir -p gpurun_out
timeout 100 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "pipelin" > gpurun_out/pytest_pipe.log 2>&1; echo "pytest pipelined rc=$?"
tail -2 gpurun_out/pytest_pipe.log
timeout 120 python -m pytest tests/test_gpu_fullsize.py -m gpu -x -q -k "drop_in_call" > gpurun_out/pytest_full.log 2>&1; echo "pytest fullsize rc=$?"
tail -2 gpurun_out/pytest_full.log
timeout 100 python bench.py --no-cpu > gpurun_out/bench_g.json 2> gpurun_out/bench_g.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_g.json").read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"], d["e2e"].get("sweep_ms"), d["clocks"])
PY
