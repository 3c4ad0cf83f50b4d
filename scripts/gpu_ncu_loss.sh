#!/bin/bash
TAG=${1:-loss}
mkdir -p gpurun_out
timeout 300 python scripts/bench_loss.py --iters 5 > /dev/null 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:loss_ -s 6 -c 3 -o gpurun_out/ncu_$TAG -f python scripts/bench_loss.py --iters 5 > gpurun_out/ncu_$TAG.log 2>&1; echo "ncu rc=$?"
tail -2 gpurun_out/ncu_$TAG.log
