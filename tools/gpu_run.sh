#!/bin/bash
# One runner for every GPU-box pass:  gpurun [--gpus N] -- bash tools/gpu_run.sh <task> [args] [-- <task> ...]
# Outputs land in gpurun_out/ (scratch); copy what should be judged into profiles/.
#   suite [pytest -k expr]   the -m gpu test suite (or a selection)
#   smoke                    __graft_entry__.smoke()
#   bench [tag] [args]       python bench.py [args]            -> bench_<tag>.json
#   ref                      bench.py --impl reference         -> bench_ref.json
#   torchbench N [tag]       bench.py under torchrun on N GPUs -> bench_<tag>.json
#   probe <cfgs...>          tools/gpu_probe.py (c1 c2 c3 c4 rad line)
#   variant <tag> <cfgs...>  tools/variant_probe.py
#   launches [tag]           ncu launch list of the bench command -> launches_<tag>.csv
#   ncu <tag> <cfg> [k]      ncu --set full of one config's kernel, summarised on the box
#   cli                      stencil_main / run.sh / run_expr.py
#   e2e                      pipelined host path tests + tools/e2e_probe.py over parallelogram counts
#   multi N                  device-group tests, multi_check and config5 under torchrun
#   n8                       the 8-GPU pass (fabric probe, bench, CLI on config 5, group tests)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
task() {
  local t=$1; shift
  case $t in
    suite) timeout 1200 python -m pytest tests -m gpu -x -q ${1:+-k "$1"} > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log ;;
    smoke) timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log; tail -8 gpurun_out/smoke.log ;;
    bench) local tag=${1:-a}; shift; timeout 900 python bench.py "$@" > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; echo "bench rc=$?"; tail -c 400 gpurun_out/bench_$tag.json ;;
    ref) timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2>&1; echo "ref rc=$?"; tail -c 300 gpurun_out/bench_ref.json ;;
    torchbench) local n=$1 tag=${2:-n$1}; timeout 900 $TR --nproc-per-node $n --master-port 295$n bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; echo "bench N=$n rc=$?" ;;
    probe) timeout 900 python tools/gpu_probe.py "$@" 2>&1 | tee gpurun_out/probe.log ;;
    variant) local tag=$1; shift; timeout 600 python tools/variant_probe.py $tag "$@" 2>&1 | grep -v Warn | tee -a gpurun_out/variants.log ;;
    launches) local tag=${1:-a}; local CMD="python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e --no-config5"
      $CMD > gpurun_out/plain_$tag.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
        --log-file gpurun_out/launches_$tag.csv $CMD > gpurun_out/ncu_list_$tag.log 2>&1; echo "launch list rc=$?" ;;
    ncu) local tag=$1; shift; local CMD="python tools/one_kernel.py $*"
      $CMD > gpurun_out/plain_$tag.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:sweep_ -s ${SKIP:-8} -c ${COUNT:-1} \
        -o gpurun_out/prof_$tag -f $CMD > gpurun_out/ncu_full_$tag.log 2>&1
      if [ $? -eq 0 ]; then { echo "# ncu --set full: tools/one_kernel.py $* (tag $tag)"; echo; python tools/ncu_summary.py gpurun_out/prof_$tag.ncu-rep;
        echo; echo "## SASS instruction mix (share of executed warp instructions)"; echo; echo '```'; python tools/ncu_mix.py gpurun_out/prof_$tag.ncu-rep 20; echo '```'; } > gpurun_out/ncu_$tag.md
        ncu -i gpurun_out/prof_$tag.ncu-rep --page raw --csv > gpurun_out/ncu_${tag}_raw.csv; fi; echo "ncu $tag done" ;;
    cli) local BIN=./stencil_b200/host/build/bin/stencil_main
      { $BIN -s 4096 -i 100 -b 64 -r 1 --dtype f64 -R 3 -m CPUOMP CUDA -c; $BIN -s 4096 -i 100 -b 64 -r 1 -R 3 -m CUDA -c; bash run.sh -R 3 -c;
        $BIN -s 512 --shape 512,512,512 --dtype f64 -i 100 -b 64 -R 3 -m CUDA; } > gpurun_out/cli.log 2>&1; tail -30 gpurun_out/cli.log
      python run_expr.py --iters 1000 100000 --repeat 2 --out gpurun_out/run_expr.csv > gpurun_out/run_expr.log 2>&1; cat gpurun_out/run_expr.csv
      python run_expr.py --size 4096 --block-as-tile --blocks 8 32 128 512 --iters 100 --repeat 2 --out gpurun_out/run_expr_tile.csv > gpurun_out/run_expr_tile.log 2>&1; cat gpurun_out/run_expr_tile.csv ;;
    e2e) timeout 600 python -m pytest tests -m gpu -x -q -k "pipelined or dropin or host" > gpurun_out/pytest_e2e.log 2>&1; tail -2 gpurun_out/pytest_e2e.log
      for c in ${CHUNKS:-auto 7 9 11 13}; do if [ $c = auto ]; then timeout 120 python tools/e2e_probe.py 2>&1 | tail -1; else SB200_PIPELINE_CHUNKS=$c timeout 120 python tools/e2e_probe.py 2>&1 | tail -1; fi; done | tee gpurun_out/e2e_probe.log ;;
    multi) local n=${1:-2}; timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -3
      timeout 300 $TR --nproc-per-node $n --master-port 29611 tools/multi_check.py 2>&1 | tail -8 | tee gpurun_out/multi_check_n$n.log
      timeout 300 $TR --nproc-per-node $n --master-port 29612 tools/config5.py --mode strong --steps 40 2>&1 | grep '^{' | tee gpurun_out/config5_n$n.log
      for ex in p2p nccl; do timeout 300 $TR --nproc-per-node $n --master-port 29613 tools/config4.py --exchange $ex 2>&1 | grep '^{'; done | tee gpurun_out/config4_n$n.log
      timeout 300 python tools/config4.py 2>&1 | grep '^{' | tee -a gpurun_out/config4_n$n.log ;;
    n8) nvidia-smi topo -m > gpurun_out/topo_n8.txt 2>&1; lscpu | grep -E "NUMA|Socket|Model name|^CPU\(s\)" > gpurun_out/lscpu_n8.txt 2>&1
      [ -x tools/pcie_probe ] || nvcc -O2 -std=c++17 -gencode arch=compute_100a,code=sm_100a -cudart static -o tools/pcie_probe tools/pcie_probe.cu
      timeout 300 tools/pcie_probe 512 > gpurun_out/pcie_probe_n8.log 2>&1; echo "probe rc=$?"
      task torchbench 8 n8
      timeout 600 stencil_b200/host/build/bin/stencil_main -s 1 -i 8 -b 1 --shape 1024,1024,2048 --dtype f64 --init random --gpus 8 -R 2 -m CUDA -c > gpurun_out/cli_config5_n8.log 2>&1; echo "cli rc=$?"
      timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/pytest_multi_n8.log 2>&1; echo "pytest rc=$?"
      tail -3 gpurun_out/pytest_multi_n8.log; tail -8 gpurun_out/cli_config5_n8.log; cat gpurun_out/pcie_probe_n8.log ;;
    *) echo "unknown task $t"; return 2 ;;
  esac
}
args=()
for a in "$@"; do if [ "$a" = "--" ]; then task "${args[@]}"; args=(); else args+=("$a"); fi; done
[ ${#args[@]} -gt 0 ] && task "${args[@]}"
