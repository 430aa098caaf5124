#!/bin/bash
# Random-scene traversal fuzz through the shipped library on the GPU (cooperative trace kernel, three builders).
mkdir -p gpurun_out
S=${1:-250}; R=${2:-2000}; SEED=${3:-7}; TAG=${4:-r02_t}
timeout 400 python tools/fuzz_traversal.py --gpu --scenes $S --rays $R --seed $SEED > gpurun_out/${TAG}_fuzz_gpu.txt 2>&1; echo rc=$? >> gpurun_out/${TAG}_fuzz_gpu.txt
tail -3 gpurun_out/${TAG}_fuzz_gpu.txt
timeout 100 python tools/fuzz_traversal.py --gpu --extreme --scenes 100 --rays 2000 --seed 3 > gpurun_out/${TAG}_fuzz_gpu_extreme.txt 2>&1; echo rc=$? >> gpurun_out/${TAG}_fuzz_gpu_extreme.txt
tail -3 gpurun_out/${TAG}_fuzz_gpu_extreme.txt
