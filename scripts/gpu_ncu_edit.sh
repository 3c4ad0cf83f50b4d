#!/bin/bash
# ncu of the texture_edit entry point: launch list + --set full of texture_edit_kernel.  Usage: bash scripts/gpu_ncu_edit.sh <tag>
TAG=${1:-r1}; OUT=gpurun_out; mkdir -p $OUT
python scripts/bench_edit.py --no-cpu --iters 5 > $OUT/bench_edit_$TAG.json 2> $OUT/bench_edit_$TAG.err; echo "bench_edit rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_edit_$TAG.csv \
  python scripts/bench_edit.py --no-cpu --iters 2 > $OUT/ncu_launch_edit_$TAG.log 2>&1; echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'texture_edit_kernel' -s 3 -c 2 \
  -o $OUT/edit_$TAG -f python scripts/bench_edit.py --no-cpu --iters 2 > $OUT/ncu_full_edit_$TAG.log 2>&1; echo "ncu full rc=$?"
tail -2 $OUT/ncu_full_edit_$TAG.log
