#!/bin/bash
# Final measurements, part B: bench lines (roofline from this build's ncu counters in profiles/traffic.json), reference arm, time to spp.
TAG=r02_i
mkdir -p gpurun_out
python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err
python bench.py --no-cpu-baseline --steps 10 --spp-target 1024 > gpurun_out/${TAG}_bench_C2_time_to_1024spp.json 2>/dev/null
python bench.py --no-cpu-baseline --steps 10 --spp-target 1024 --spp-criterion min > gpurun_out/${TAG}_bench_C2_time_to_1024spp_min.json 2>/dev/null
for w in C1 C3 C4; do
  python bench.py --workload $w --steps 10 > gpurun_out/${TAG}_bench_$w.json 2> gpurun_out/${TAG}_bench_$w.err
done
python bench.py --workload C5 --steps 10 --spp-target 16384 > gpurun_out/${TAG}_bench_C5.json 2> gpurun_out/${TAG}_bench_C5.err
for f in gpurun_out/${TAG}_bench*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    cb = d.get("cpu_baseline") or {}
    rf = d.get("roofline") or {}
    print(sys.argv[1].split("/")[-1], round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "paths", round(d.get("mpaths_per_s", 0), 1), "kernel_ms", d.get("kernel_ms"),
          "roofline", rf.get("bound"), rf.get("frac"), "traffic", rf.get("traffic"), "cpu", cb.get("value"), cb.get("cores"), "spp", d.get("time_to_spp"))
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
