#!/bin/bash
# A/B of build_exp/libptb_<tag>.so builds (see _build.build_experiment): bench.py per library, repeated.
mkdir -p gpurun_out
out=gpurun_out/${OUT:-ab_run.txt}
: > $out
for rep in 1 2; do
for lib in ${LIBS:-default}; do
  if [ "$lib" = default ]; then unset PTB_LIB; else export PTB_LIB=$PWD/build_exp/libptb_$lib.so; fi
  python bench.py --workload ${WORKLOAD:-C2} --steps ${STEPS:-20} --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 | \
    python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$lib', round(d['value'],1), 'trace', round(d['kernel_ms']['trace'],4), 'scatter', round(d['kernel_ms']['scatter'],4), 'e2e', round(d['e2e']['value'],1))" >> $out 2>&1
done; done
cat $out
