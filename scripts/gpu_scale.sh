#!/bin/bash
# Multi-GPU bench line (launched the way the driver does).  Usage: bash scripts/gpu_scale.sh <N> <tag>
N=${1:-2}; TAG=${2:-scale}
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus $N --steps 5 --warmup 3 > $OUT/bench_${TAG}_n$N.json 2> $OUT/bench_${TAG}_n$N.err; echo "bench N=$N rc=$?"
tail -3 $OUT/bench_${TAG}_n$N.err
cat $OUT/bench_${TAG}_n$N.json
