#!/bin/bash
# Random-material scatter fuzz through the shipped library on the GPU (wavefront + megakernel) against the oracle.
mkdir -p gpurun_out
S=${1:-150}; SEED=${2:-5}; TAG=${3:-r02_w}
timeout ${4:-150} python tools/fuzz_scatter.py --gpu --scenes $S --seed $SEED --width 48 --height 32 > gpurun_out/${TAG}_fuzz_scatter_gpu.txt 2>&1; echo rc=$? >> gpurun_out/${TAG}_fuzz_scatter_gpu.txt
tail -4 gpurun_out/${TAG}_fuzz_scatter_gpu.txt
