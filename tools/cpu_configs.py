"""CPU baseline of every BASELINE config on the GPU box's host cores: the oracle port
(reference-order naive loop, src/stencil/stencil.cpp:94-131 generalised) with 1 thread and
with all threads, a bounded number of steps each (throughput does not depend on T).
    python tools/cpu_configs.py > gpurun_out/cpu_configs.log"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle

CASES = [
    ("1: 2-D 5-pt 4096^2 fp64", (4096, 4096), oracle.STAR, oracle.F64),
    ("1: 2-D 5-pt 4096^2 fp32 (reference dtype)", (4096, 4096), oracle.STAR, oracle.F32),
    ("2: 3-D 7-pt 512^3 fp64", (512, 512, 512), oracle.STAR, oracle.F64),
    ("3: 3-D 27-pt 512^3 fp64", (512, 512, 512), oracle.BOX, oracle.F64),
    ("4: 2-D 9-pt 16384^2 fp32", (16384, 16384), oracle.BOX, oracle.F32),
    ("5: 3-D 7-pt 1024x1024x256 fp64 (one of 8 slabs of 2048x1024x1024)", (1024, 1024, 256), oracle.STAR, oracle.F64),
]
nthr = oracle.max_threads()
print(f"host threads available: {nthr}")
for name, n, shape, dt in CASES:
    a = oracle.fill(n, 1, dt, oracle.INIT_RANDOM, seed=1)
    b = oracle.fill(n, 1, dt, oracle.INIT_REF)
    pts = 1
    for v in n:
        pts *= v
    for threads in (1, nthr):
        t0 = time.perf_counter()
        oracle.sweep(n, 1, shape, 1, a, b, threads)
        one = time.perf_counter() - t0
        steps = max(1, min(50, int(4.0 / max(one, 1e-4))))
        t0 = time.perf_counter()
        oracle.sweep(n, 1, shape, steps, a, b, threads)
        dt_s = time.perf_counter() - t0
        print(f"{name:70s} threads={threads:3d} steps={steps:3d}  {pts * steps / dt_s / 1e9:8.3f} GStencil/s", flush=True)
