#!/bin/bash
mkdir -p gpurun_out
out=gpurun_out/r02_k_ab_class_split.txt; : > $out
for lib in default split; do
  if [ "$lib" = default ]; then unset PTB_LIB; else export PTB_LIB=$PWD/build_exp/libptb_$lib.so; fi
  r=$(timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_baseline_configs.py -m gpu -x -q -k "queues_equal or reference_shader_vectors or first_bounce or C1 or C3" 2>&1 | tail -1)
  echo "$lib tests: $r" >> $out
done
for rep in 1 2; do for lib in default split; do
  if [ "$lib" = default ]; then unset PTB_LIB; else export PTB_LIB=$PWD/build_exp/libptb_$lib.so; fi
  for w in C3 C1 C4; do
  python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$lib $w', round(d['value'],1), 'trace', round(d['kernel_ms']['trace'],4), 'scatter', round(d['kernel_ms']['scatter'],4))" >> $out 2>&1
  done
done; done
cat $out
