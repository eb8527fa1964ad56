#!/bin/bash
# Round-end validation on one GPU: full GPU test suite, the default bench line, the ncu launch list and one full capture of the
# dominant kernel (tools/profile_bench.sh recipe), all into gpurun_out/<tag>_*.
TAG=${1:-r02final}
mkdir -p gpurun_out
timeout 700 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest_gpu.log
timeout 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; tail -c 400 gpurun_out/${TAG}_bench.err
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-legs --no-spgemm"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1; echo "launch list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_spmv_red -s 8 -c 2 -o gpurun_out/${TAG}_spmv_red -f $CMD > gpurun_out/${TAG}_ncu2.log 2>&1; echo "capture rc=$?"
ncu -i gpurun_out/${TAG}_spmv_red.ncu-rep --page raw --csv --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,launch__registers_per_thread,sm__warps_active.avg.pct_of_peak_sustained_active,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,lts__t_sector_hit_rate.pct > gpurun_out/${TAG}_spmv_red_metrics.csv 2>&1
tail -3 gpurun_out/${TAG}_spmv_red_metrics.csv | cut -c1-400
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/${TAG}_bench.json") if l.startswith("{")][-1])
print({k:d.get(k) for k in ("value","ms_per_step","gpu_launches")}, d["roofline"]["frac"], d["roofline"]["avg_launch_ms"], d["roofline"]["kernel_share_of_step"])
print("e2e", d.get("e2e")); print("cpu", d.get("cpu_baseline"))
for k in ("spgemm","bfs","sssp","tc"):
    v=d.get(k) or {}
    print(k, {x:v.get(x) for x in ("value","unit","ms","ms_per_run","error","parity_bit_exact")})
PY
