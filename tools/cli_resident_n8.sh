BIN=stencil_b200/host/build/bin/stencil_main
{ echo '$ stencil_main --shape 1024,1024,2048 --dtype f64 -i 40 --init random --resident --gpus 8 -R 3 -m CUDA'
$BIN -s 1 -b 1 --shape 1024,1024,2048 --dtype f64 -i 40 --init random --resident --gpus 8 -R 3 -m CUDA
echo '$ same, --gpus 1'
$BIN -s 1 -b 1 --shape 1024,1024,2048 --dtype f64 -i 40 --init random --resident --gpus 1 -R 2 -m CUDA
echo '$ stencil_main --shape 1024,1024,256 --dtype f64 -i 8 --init random --resident --gpus 8 -m CUDA -c'
$BIN -s 1 -b 1 --shape 1024,1024,256 --dtype f64 -i 8 --init random --resident --gpus 8 -m CUDA -c
echo '$ stencil_main -s 16384 --stencil box -i 48 --init random --resident --gpus 8 -R 3 -m CUDA   (config 4 grid, fp32)'
$BIN -s 16384 -b 1 --stencil box -i 48 --init random --resident --gpus 8 -R 3 -m CUDA
echo '$ same, --gpus 1'
$BIN -s 16384 -b 1 --stencil box -i 48 --init random --resident --gpus 1 -R 2 -m CUDA
echo '$ stencil_main -s 8192 --stencil box -i 20 --init random --resident --gpus 8 -m CUDA -c'
$BIN -s 8192 -b 1 --stencil box -i 20 --init random --resident --gpus 8 -m CUDA -c; } > gpurun_out/cli_resident_n8.log 2>&1
python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -2 >> gpurun_out/cli_resident_n8.log
cat gpurun_out/cli_resident_n8.log
