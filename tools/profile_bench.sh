#!/bin/bash
# Profiling recipe of /opt/skills/guides/B200_PROFILING.md for the bench workload (run under gpurun, 1 GPU).
# Usage: tools/profile_bench.sh <tag>   -> gpurun_out/<tag>_{plain.log,launches.csv,prof.ncu-rep}
set -u
TAG=${1:-r01}
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline ${BENCH_ARGS:-}"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:${KREGEX:-k_spmv_stream} -s 8 -c 2 -o gpurun_out/${TAG}_prof -f $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
ls -la gpurun_out/
