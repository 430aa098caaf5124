#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 tools/reduce_ab.py > gpurun_out/r02_m_reduce_ab_${N}gpu.json 2> gpurun_out/r02_m_reduce_ab_${N}gpu.err
tail -1 gpurun_out/r02_m_reduce_ab_${N}gpu.json | python -c "
import sys, json
d = json.loads(sys.stdin.read())
for size, v in d['sizes'].items():
    for m, r in v['modes'].items(): print(d['n_gpus'], size, m, round(r['exchange_resolve_ms_device'], 4), 'ms device;', round(r['call_ms_wall_on_this_rank'], 3), 'ms call;', r['checksum'])
" || tail -5 gpurun_out/r02_m_reduce_ab_${N}gpu.err
