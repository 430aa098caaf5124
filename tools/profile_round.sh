#!/bin/bash
# One gpurun call: launch lists and full ncu captures of the current build (see /opt/skills/guides/B200_PROFILING.md).
# usage (on the GPU box): bash tools/profile_round.sh r01_v9
TAG=${1:-prof}
mkdir -p gpurun_out
BENCH="python bench.py --no-cpu-baseline --steps 2 --warmup 3 --frames 4"
BUILD="python tools/build_probe.py"
$BENCH > gpurun_out/${TAG}_plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $BENCH > gpurun_out/${TAG}_ncu_launches.log 2>&1
$BENCH > gpurun_out/${TAG}_plain_bench2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"k_trace_coop|k_scatter" -s 24 -c 2 -f -o gpurun_out/${TAG} $BENCH > gpurun_out/${TAG}_ncu_full.log 2>&1
$BUILD > gpurun_out/${TAG}_plain_build.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 1200 --csv --log-file gpurun_out/${TAG}_builder_launches.csv $BUILD > gpurun_out/${TAG}_ncu_build.log 2>&1
ls -la gpurun_out | tail -20
