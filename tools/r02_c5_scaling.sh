#!/bin/bash
# BASELINE config 5 as written: 4K, 10M triangles, 16384 spp sample-split over N GPUs, exchange + resolve inside the timed region.
# usage (on an N-GPU box): bash tools/r02_c5_scaling.sh N [SPP] [extra: tests]
N=${1:-2}; SPP=${2:-16384}
mkdir -p gpurun_out
if [ "$3" = tests ]; then
  (time timeout 900 python -m pytest tests/test_multi_gpu.py tests/test_headless_app.py -m gpu -x -q --durations=6) > gpurun_out/r02_e_multigpu_tests_${N}gpu.log 2>&1
  tail -4 gpurun_out/r02_e_multigpu_tests_${N}gpu.log
fi
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N \
  --workload C5 --no-cpu-baseline --steps 5 --warmup 3 --spp-target $SPP $SPPARGS > gpurun_out/r02_e_bench_C5_${N}gpu_spp${SPP}.json 2> gpurun_out/r02_e_bench_C5_${N}gpu.err
python - <<PY
import json
try:
    d = json.load(open("gpurun_out/r02_e_bench_C5_${N}gpu_spp${SPP}.json"))
    print("C5 N=$N", round(d["value"], 1), "Mrays/s", "e2e", round(d["e2e"]["value"], 1), d["time_to_spp"], d.get("setup_seconds"))
except Exception as e:
    print("unreadable", e); print(open("gpurun_out/r02_e_bench_C5_${N}gpu.err").read()[-2000:])
PY
