#!/bin/bash
# Bench lines of the final build (after the per-face attribute change); time-to-spp lines of r02_i stay (build before it: 1 % slower).
TAG=r02_r
mkdir -p gpurun_out
python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err
python bench.py --no-cpu-baseline --steps 10 --spp-target 1024 > gpurun_out/${TAG}_bench_C2_time_to_1024spp.json 2>/dev/null
for w in C1 C3 C4 C5; do
  python bench.py --workload $w --steps 10 > gpurun_out/${TAG}_bench_$w.json 2> gpurun_out/${TAG}_bench_$w.err
done
for f in gpurun_out/${TAG}_bench*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    cb = d.get("cpu_baseline") or {}
    rf = d.get("roofline") or {}
    rs = d.get("roofline_scatter") or {}
    print(sys.argv[1].split("/")[-1], round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), round((d["e2e"].get("frames_per_step_1") or {}).get("value", 0), 1), "paths", round(d.get("mpaths_per_s", 0), 1),
          "ms", d.get("kernel_ms"), "issue", rf.get("frac"), rf.get("lanes_per_instruction"), rf.get("traffic"), "scatter", rs.get("frac"), rs.get("traffic"), "cpu", cb.get("value"), cb.get("cores"), cb.get("kind"), "spp", (d.get("time_to_spp") or {}).get("seconds"))
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
