#!/bin/bash
# ncu capture of k_pull_any in variants A (in place) and B (second vector): tools/debug/bfs_variants.py prints A, B, C, A, B twice each
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_pull_any -c 12 -o gpurun_out/pull_any_variants -f \
  python tools/debug/bfs_variants.py 24 AB > gpurun_out/pull_any_variants.log 2>&1
echo rc=$?
ncu -i gpurun_out/pull_any_variants.ncu-rep --page raw --csv --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,lts__t_sector_hit_rate.pct,l1tex__t_bytes_pipe_lsu_mem_global_op_ld.sum,l1tex__t_bytes_pipe_lsu_mem_global_op_st.sum,smsp__inst_executed.sum,lts__t_sectors_srcunit_tex_op_write.sum,lts__t_sectors_srcunit_tex_op_read.sum,lts__t_sectors_srcnode_gpc_op_read.sum > gpurun_out/pull_any_variants.csv 2>&1
tail -14 gpurun_out/pull_any_variants.csv | cut -c1-600
