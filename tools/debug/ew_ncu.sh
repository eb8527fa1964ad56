#!/bin/bash
# ncu capture of the element-wise kernels of a level BFS (one GPU, scale 24): k_assign_scalar and k_vec_writeback
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_assign_scalar|k_vec_writeback" --launch-skip 14 -c 14 -o gpurun_out/ew_kernels -f \
  python tools/debug/bfs_variants.py 24 A > gpurun_out/ew_kernels.log 2>&1
echo rc=$?
ncu -i gpurun_out/ew_kernels.ncu-rep --page raw --csv --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio,smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,launch__grid_size > gpurun_out/ew_kernels.csv 2>&1
cut -d, -f5,9,12- gpurun_out/ew_kernels.csv | cut -c1-260 | tail -16
