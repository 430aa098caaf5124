#!/bin/bash
# GPU run 3 of round 2: tests after the host-math change, ray-order experiment, like-for-like integrator rows.
mkdir -p gpurun_out
(time timeout 900 python -m pytest tests -m gpu -x -q) > gpurun_out/r02_d_gputests.log 2>&1
tail -3 gpurun_out/r02_d_gputests.log
for w in C2 C3; do
  timeout 600 python tools/ray_order_experiment.py --workload $w --out gpurun_out/r02_d_ray_order_$w.json > gpurun_out/r02_d_ray_order_$w.log 2>&1
  tail -2 gpurun_out/r02_d_ray_order_$w.log | cut -c1-1500
done
B="python bench.py --no-cpu-baseline --steps 10 --warmup 3"
$B --trace-variant 0 > gpurun_out/r02_d_bench_variant0.json 2> gpurun_out/r02_d_bench_variant0.err
$B --integrator megakernel > gpurun_out/r02_d_bench_megakernel.json 2> gpurun_out/r02_d_bench_megakernel.err
for f in variant0 megakernel; do python -c "
import json,sys
d=json.load(open('gpurun_out/r02_d_bench_$f.json')); print('$f', round(d['value'],1), d['kernel_ms'])"; done
