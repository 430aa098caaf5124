#!/bin/bash
# 8-GPU box: multi-GPU tests at 2/4/8 ranks, C5 as written (16384 spp, mean criterion), C2 headline at 8 GPUs.
mkdir -p gpurun_out
bash tools/r02_c5_scaling.sh 8 16384 tests
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus 8 --no-cpu-baseline --spp-target 1024 \
  > gpurun_out/r02_e_bench_C2_8gpu.json 2> gpurun_out/r02_e_bench_C2_8gpu.err
python -c "
import json
d=json.load(open('gpurun_out/r02_e_bench_C2_8gpu.json')); print('C2 N=8', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d['e2e'].get('exchange'), d['time_to_spp'], d.get('setup_seconds'))"
timeout 300 path_tracer_b200/bin/ptb_headless --scene spherefield:488:5:0 --size 3840x2160 --spp 256 --devices 0,1,2,3,4,5,6,7 --table path_tracer_b200/data/sRGBSpectrumTable.dat --out gpurun_out/r02_e_c5_8gpu_256spp.png > gpurun_out/r02_e_headless_c5_8gpu.json 2> gpurun_out/r02_e_headless_c5_8gpu.err
cat gpurun_out/r02_e_headless_c5_8gpu.json; tail -2 gpurun_out/r02_e_headless_c5_8gpu.err
ls -la gpurun_out/r02_e_c5_8gpu_256spp.png
