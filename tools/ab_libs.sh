#!/bin/bash
# A/B of alternative in-tree builds (build_exp/libptb_*.so) on the GPU box: bench.py per library and variant.
out=${1:-gpurun_out/ab_libs.txt}
: > "$out"
for lib in ${LIBS:-default}; do
  for v in ${VARIANTS:-10}; do
    if [ "$lib" = default ]; then unset PTB_LIB; else export PTB_LIB=$PWD/build_exp/libptb_$lib.so; fi
    python bench.py --steps ${STEPS:-10} --warmup 3 --no-cpu-baseline --trace-variant $v 2>/dev/null | tail -1 | \
      python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$lib variant $v', round(d['value'],1), 'trace', round(d['kernel_ms']['trace'],4), 'scatter', round(d['kernel_ms']['scatter'],4))" >> "$out" 2>&1
  done
done
cat "$out"
