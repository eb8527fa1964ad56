#!/bin/bash
# N=<gpus> SCALES="22 24" bash tools/debug/run_bfs.sh : level BFS sharded by source block (bench.py --workload bfs)
N=${N:-2}
mkdir -p gpurun_out
for S in ${SCALES:-22 24}; do
  if [ "$N" = 1 ]; then
    timeout -k 5 170 python bench.py --workload bfs --scale $S --steps 5 $EXTRA > gpurun_out/bfs_s${S}_n1.json 2> gpurun_out/bfs_s${S}_n1.err
  else
    timeout -k 5 170 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
      bench.py --workload bfs --gpus $N --scale $S --steps 5 $EXTRA > gpurun_out/bfs_s${S}_n${N}.json 2> gpurun_out/bfs_s${S}_n${N}.err
  fi
  echo "rc=$? scale $S N=$N"; tail -c 1500 gpurun_out/bfs_s${S}_n${N}.json; tail -3 gpurun_out/bfs_s${S}_n${N}.err
done
