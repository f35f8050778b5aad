#!/bin/bash
# The reference-facing CLI on the GPU box: BASELINE configs[0] through stencil_main, run.sh,
# and a short run_expr.py sweep (reference block sizes, reduced iteration list).
mkdir -p gpurun_out
BIN=./stencil_b200/host/build/bin/stencil_main
L=gpurun_out/cli.log
{ echo "\$ stencil_main -s 4096 -i 100 -b 64 -r 1 --dtype f64 -R 3 -m CPU CPUOMP CUDA -c";
  $BIN -s 4096 -i 100 -b 64 -r 1 --dtype f64 -R 3 -m CPUOMP CUDA -c;
  echo; echo "\$ stencil_main -s 4096 -i 100 -b 64 -r 1 -R 3 -m CUDA -c   (fp32, the reference dtype)";
  $BIN -s 4096 -i 100 -b 64 -r 1 -R 3 -m CUDA -c;
  echo; echo "\$ ./run.sh -R 3 -c"; bash run.sh -R 3 -c;
  echo; echo "\$ stencil_main --shape 512,512,512 --dtype f64 -i 100 -b 64 -R 3 -m CUDA";
  $BIN -s 512 --shape 512,512,512 --dtype f64 -i 100 -b 64 -R 3 -m CUDA; } > $L 2>&1
tail -40 $L
python run_expr.py --iters 1000 100000 --repeat 2 --out gpurun_out/run_expr.csv > gpurun_out/run_expr.log 2>&1; cat gpurun_out/run_expr.csv
