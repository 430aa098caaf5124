#!/bin/bash
# After the host-builder rewrite (threaded gather / ranks / collapse / emit): smoke, the GPU suite, and the host build times
# of C2 and C4 on the GPU box's cores (bench lines carry wide_bvh.build_seconds; PTB_WIDE_TIMING prints the phases).
TAG=r02_s
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
PTB_WIDE_TIMING=1 python bench.py --no-cpu-baseline --steps 3 --warmup 3 > gpurun_out/${TAG}_bench_C2.json 2> gpurun_out/${TAG}_bench_C2.err
PTB_WIDE_TIMING=1 python bench.py --workload C4 --no-cpu-baseline --steps 3 --warmup 3 > gpurun_out/${TAG}_bench_C4.json 2> gpurun_out/${TAG}_bench_C4.err
nproc
for w in C2 C4; do grep "ptb wide build" gpurun_out/${TAG}_bench_$w.err | tr '\n' ';'; echo; python - gpurun_out/${TAG}_bench_$w.json <<'PY'
import json, sys
d = json.load(open(sys.argv[1]))
print(sys.argv[1].split("/")[-1], round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "ms", d.get("kernel_ms"), "wide", d["config"].get("wide_bvh"))
PY
done
