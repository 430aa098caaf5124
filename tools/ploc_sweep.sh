mkdir -p gpurun_out
for w in C2 C4; do for r in 8 16 32 64 128; do
  PTB_PLOC_RADIUS=$r timeout 600 python bench.py --workload $w --builder ploc-gpu --no-cpu-baseline --steps 8 > gpurun_out/ploc_${w}_r$r.json 2>/dev/null
  python -c "
import json;d=json.load(open('gpurun_out/ploc_${w}_r$r.json'));w=d['config']['wide_bvh'];print('$w radius $r: %.1f Mrays/s trace %.3f ms  build %.1f ms  nodes %d sah %.2f'%(d['value'],d['kernel_ms']['trace'],w['device_build_ms'],w['node_count'],w['sah_cost']))"
done; done | tee gpurun_out/ploc_radius_sweep.txt
