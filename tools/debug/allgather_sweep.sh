#!/bin/bash
N=${1:-4}
run() { env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 tools/debug/allgather_bench.py 2>/dev/null | grep world; }
run NCCL_DEBUG=WARN
run NCCL_ALGO=Ring
run NCCL_ALGO=NVLS
run NCCL_MIN_NCHANNELS=32
run NCCL_ALGO=Ring NCCL_MIN_NCHANNELS=32
run NCCL_PROTO=Simple NCCL_MIN_NCHANNELS=32
run NCCL_NVLS_ENABLE=0
