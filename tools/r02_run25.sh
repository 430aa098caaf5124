#!/bin/bash
mkdir -p gpurun_out
out=gpurun_out/r02_q_ab_face_attrs.txt; : > $out
r=$(timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_baseline_configs.py -m gpu -x -q 2>&1 | tail -1); echo "face_attrs tests: $r" >> $out
for rep in 1 2; do for off in 0 1; do
  for w in C2 C5 C4 C3; do
  if [ $off = 1 ]; then export PTB_NO_FACE_ATTRS=1; else unset PTB_NO_FACE_ATTRS; fi
  python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('indexed' if $off else 'face_attrs', '$w', round(d['value'],1), 'trace', round(d['kernel_ms']['trace'],4), 'scatter', round(d['kernel_ms']['scatter'],4))" >> $out 2>&1
  done
done; done
cat $out
