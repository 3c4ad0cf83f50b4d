#!/bin/bash
N=${1:-2}; TAG=${2:-sweep}
mkdir -p gpurun_out
P=$((29500 + RANDOM % 2000))
PEER_SWEEP=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P scripts/peer_allreduce_bench.py 2>&1 | grep -v "^W\|^\*\*\*\|Warn" | tee gpurun_out/peer_allreduce_${TAG}_n${N}.txt | tail -40
