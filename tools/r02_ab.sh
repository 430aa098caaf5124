#!/bin/bash
# A/B of experiment builds: correctness subset first (bit-identity tests), then bench per library, twice.
mkdir -p gpurun_out
out=gpurun_out/${OUT:-r02_f_ab.txt}
: > $out
for lib in ${LIBS:-default}; do
  if [ "$lib" = default ]; then unset PTB_LIB; else export PTB_LIB=$PWD/build_exp/libptb_$lib.so; fi
  r=$(timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "raycast_wide_bvh or trace_variants or megakernel_equals or full_size_c2 or first_bounce" 2>&1 | tail -1)
  echo "$lib tests: $r" >> $out
done
for rep in 1 2; do
for lib in ${LIBS:-default}; do
  if [ "$lib" = default ]; then unset PTB_LIB; else export PTB_LIB=$PWD/build_exp/libptb_$lib.so; fi
  for w in ${WORKLOADS:-C2}; do
  python bench.py --workload $w --steps ${STEPS:-20} --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 | \
    python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$lib $w', round(d['value'],1), 'trace', round(d['kernel_ms']['trace'],4), 'scatter', round(d['kernel_ms']['scatter'],4), 'e2e', round(d['e2e']['value'],1))" >> $out 2>&1
  done
done; done
cat $out
