#!/bin/bash
N=${1:-2}; TAG=${2:-peer}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_peer_allreduce.py -x -q 2>&1 | tail -4
P=$((29500 + RANDOM % 2000))
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P scripts/peer_allreduce_bench.py 2>&1 | grep -v "^W\|^\*\*\*\|Warn" | tee gpurun_out/peer_allreduce_${TAG}_n${N}.txt | tail -12
