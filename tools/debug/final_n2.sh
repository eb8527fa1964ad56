#!/bin/bash
# The driver's N > 1 launch of the default line (both legs included), on N GPUs of one box
N=${N:-2}; TAG=${1:-r02final}
mkdir -p gpurun_out
S=$(date +%s)
timeout -k 5 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
  bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_n${N}.json 2> gpurun_out/${TAG}_bench_n${N}.err
echo "rc=$? wall=$(( $(date +%s) - S ))s"; tail -c 600 gpurun_out/${TAG}_bench_n${N}.err
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/${TAG}_bench_n${N}.json") if l.startswith("{")][-1])
print({k:d.get(k) for k in ("value","ms_per_step","gpu_launches","n_gpus")}, d["roofline"]["frac"], d.get("parity_vs_single_gpu"))
print("e2e", d.get("e2e"))
for k in ("sssp","bfs"):
    v=d.get(k) or {}
    print(k, {x:v.get(x) for x in ("value","ms","ms_per_run","error","directions","parity_vs_single_gpu","iterations")})
PY
