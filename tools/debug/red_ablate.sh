#!/bin/bash
# timing sweep of k_spmv_red over the shared-memory budget (hub table size vs L1 left for misses in flight)
for kb in 100 130 162 180 194 224; do
  echo -n "smem_kb=$kb  "
  python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --smem-kb $kb 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.readlines()[-1]); print(d['roofline']['avg_launch_ms'], d['ms_per_step'])"
done
