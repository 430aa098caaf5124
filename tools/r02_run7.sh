#!/bin/bash
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "staged or queues_equal or megakernel_equals or first_bounce or full_size_c2 or ragged" 2>&1 | tail -15) > gpurun_out/r02_g_staged_tests.log
tail -6 gpurun_out/r02_g_staged_tests.log
out=gpurun_out/r02_g_ab_scatter_staged.txt; : > $out
for rep in 1 2; do for m in auto direct-staged; do
  python bench.py --steps 20 --warmup 3 --no-cpu-baseline --scatter-mode $m 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('C2 $m', round(d['value'],1), 'trace', round(d['kernel_ms']['trace'],4), 'scatter', round(d['kernel_ms']['scatter'],4), 'e2e', round(d['e2e']['value'],1), 'scatter hbm frac', round(d['roofline_scatter']['frac'],3))" >> $out 2>&1
done; done
for m in direct direct-staged queues; do
  python bench.py --workload C3 --steps 10 --warmup 3 --no-cpu-baseline --scatter-mode $m 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('C3 $m', round(d['value'],1), 'trace', round(d['kernel_ms']['trace'],4), 'scatter', round(d['kernel_ms']['scatter'],4))" >> $out 2>&1
done
cat $out
