#!/bin/bash
# Runs bench.py once per trace-kernel variant and collects the JSON lines (GPU box).
out=${1:-gpurun_out/variant_sweep.txt}
: > "$out"
for v in ${VARIANTS:-0 1 2 3 4 5 6 7 8 9 10 11 12 13 14 15}; do
  python bench.py --steps ${STEPS:-10} --warmup 3 --no-cpu-baseline --trace-variant $v 2>/dev/null | tail -1 | \
    python -c "import sys,json; d=json.loads(sys.stdin.read()); print('variant $v', d['value'], d['unit'], 'ms/step', d['ms_per_step'], 'kernel_ms', d.get('kernel_ms'), 'e2e', d['e2e']['value'])" >> "$out" 2>&1
done
cat "$out"
