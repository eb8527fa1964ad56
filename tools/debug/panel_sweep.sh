#!/bin/bash
# k_spmv_red per-launch time and step time against the column-panel threshold (bench.py --panel-kb)
N=${1:-1}
for kb in ${KBS:-0 4096 8192 12288 24576}; do
  echo -n "gpus=$N panel_kb=$kb  "
  if [ "$N" = 1 ]; then
    python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --panel-kb $kb 2>/dev/null
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus $N --steps 10 --warmup 3 --no-e2e --panel-kb $kb 2>/dev/null
  fi | python -c "import json,sys; d=json.loads(sys.stdin.readlines()[-1]); print(d['value'], d['roofline']['avg_launch_ms'], d['roofline']['launches_timed'], d['ms_per_step'])"
done
