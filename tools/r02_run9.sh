#!/bin/bash
mkdir -p gpurun_out
out=gpurun_out/r02_g_ab_scatter_staged2.txt; : > $out
for lib in default carve100 carve90; do
  if [ "$lib" = default ]; then unset PTB_LIB; else export PTB_LIB=$PWD/build_exp/libptb_$lib.so; fi
  PTB_DEBUG_OCCUPANCY=1 python bench.py --steps 20 --warmup 3 --no-cpu-baseline 2> gpurun_out/err_$lib.txt | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$lib C2 staged', round(d['value'],1), 'trace', round(d['kernel_ms']['trace'],4), 'scatter', round(d['kernel_ms']['scatter'],4))" >> $out 2>&1
  grep k_scatter_staged gpurun_out/err_$lib.txt | head -2 >> $out
done
cat $out
