#!/bin/bash
# Final measurements, part A: the whole GPU suite, then one ncu --set full capture per workload (after a plain run of the same command).
mkdir -p gpurun_out
(time timeout 1200 python -m pytest tests -m gpu -x -q --durations=8) > gpurun_out/r02_h_gputests.log 2>&1
tail -14 gpurun_out/r02_h_gputests.log
for w in C2 C1 C3 C4 C5; do
  B="python bench.py --workload $w --no-cpu-baseline --steps 2 --warmup 3 --frames 4"
  $B > gpurun_out/r02_h_plain_$w.log 2>&1 || { echo "plain run of $w failed"; tail -5 gpurun_out/r02_h_plain_$w.log; continue; }
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_trace_coop|k_scatter|k_classify" -s 30 -c 8 -f -o gpurun_out/r02_h_$w $B > gpurun_out/r02_h_ncu_$w.log 2>&1
  ls -la gpurun_out/r02_h_$w.ncu-rep 2>&1 | tail -1
done
B="python bench.py --no-cpu-baseline --steps 2 --warmup 3 --frames 4"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_h_launches.csv $B > gpurun_out/r02_h_ncu_launches.log 2>&1
