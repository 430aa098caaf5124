#!/bin/bash
# Final measurements of a round on one GPU: every workload, time to 1024 spp, the reference arm.
TAG=${1:-final}
mkdir -p gpurun_out
python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err
python bench.py --no-cpu-baseline --steps 10 --spp-target 1024 > gpurun_out/${TAG}_bench_C2_time_to_1024spp.json 2>/dev/null
for w in C1 C3 C4 C5; do
  python bench.py --workload $w --steps 10 > gpurun_out/${TAG}_bench_$w.json 2> gpurun_out/${TAG}_bench_$w.err
done
for b in lbvh-gpu ploc-gpu; do for w in C2 C4; do
  python bench.py --workload $w --builder $b --no-cpu-baseline --steps 10 > gpurun_out/${TAG}_bench_${w}_${b/-/_}_builder.json 2>/dev/null
done; done
for f in gpurun_out/${TAG}_bench*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    cb = d.get("cpu_baseline") or {}
    print(sys.argv[1].split("/")[-1], round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "paths", round(d.get("mpaths_per_s", 0), 1), "kernel_ms", d.get("kernel_ms"),
          "cpu", cb.get("value"), cb.get("cores"), "wide", (d.get("config") or {}).get("wide_bvh"), "spp", d.get("time_to_spp"))
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
