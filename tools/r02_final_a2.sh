#!/bin/bash
# Final measurements, part A again without the test suite: reports stay in /tmp on the box (together they exceed what gpurun
# brings back); their raw-page CSVs, summaries and the C2 report come back.
mkdir -p gpurun_out
for w in C2 C1 C3 C4 C5; do
  B="python bench.py --workload $w --no-cpu-baseline --steps 2 --warmup 3 --frames 4"
  $B > gpurun_out/r02_h_plain_$w.log 2>&1 || { echo "plain run of $w failed"; tail -5 gpurun_out/r02_h_plain_$w.log; continue; }
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_trace_coop|k_scatter|k_classify" -s 30 -c 8 -f -o /tmp/r02_h_$w $B > gpurun_out/r02_h_ncu_$w.log 2>&1
  ncu -i /tmp/r02_h_$w.ncu-rep --page raw --csv --print-units base > gpurun_out/r02_h_${w}_raw.csv 2>/dev/null
  python tools/ncu_summary.py /tmp/r02_h_$w.ncu-rep > gpurun_out/r02_h_${w}_ncu_summary.txt 2>/dev/null
  python tools/ncu_lines.py /tmp/r02_h_$w.ncu-rep k_trace_coop 60 > gpurun_out/r02_h_${w}_trace_source_lines.txt 2>/dev/null
  ls -la gpurun_out/r02_h_${w}_raw.csv | tail -1
done
python tools/ncu_lines.py /tmp/r02_h_C2.ncu-rep "k_scatter" 60 > gpurun_out/r02_h_C2_scatter_source_lines.txt 2>/dev/null
cp /tmp/r02_h_C2.ncu-rep gpurun_out/
B="python bench.py --no-cpu-baseline --steps 2 --warmup 3 --frames 4"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_h_launches.csv $B > gpurun_out/r02_h_ncu_launches.log 2>&1
du -sh gpurun_out
