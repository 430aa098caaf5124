#!/bin/bash
# BVH8 builder knobs (leaf size of the binary build, primitive cost of the SAH collapse) against trace time (GPU box).
out=${1:-gpurun_out/build_sweep.txt}
: > "$out"
for ml in ${MAXLEAFS:-1 2 3 4}; do
  for cp in ${CPRIMS:-0.4 1.0 1.6 2.5}; do
    PTB_WIDE_MAXLEAF=$ml PTB_WIDE_CPRIM=$cp python bench.py --steps ${STEPS:-10} --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 | \
      python -c "import sys,json; d=json.loads(sys.stdin.read()); print('maxleaf $ml cprim $cp', round(d['value'],1), 'trace', round(d['kernel_ms']['trace'],4), 'nodes', d['config']['wide_bvh']['node_count'])" >> "$out" 2>&1
  done
done
cat "$out"
