#!/bin/bash
# One GPU-box pass: parity tests, bench line, ncu launch list, ncu --set full of the compositing and binning kernels.
# Usage (from the repo root, under gpurun): bash scripts/gpu_round.sh <tag>
TAG=${1:-r1}
OUT=gpurun_out
mkdir -p $OUT
python -m pytest tests -m gpu -x -q > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_$TAG.log
tail -5 $OUT/pytest_$TAG.log
python bench.py --steps 5 --warmup 3 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench rc=$?"
cat $OUT/bench_$TAG.json
python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_ref_$TAG.json 2>> $OUT/bench_$TAG.err; echo "ref rc=$?"
cat $OUT/bench_ref_$TAG.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/launches_$TAG.csv \
  python bench.py --steps 1 --warmup 1 --views-per-gpu 2 --no-ref-cuda --no-cpu-baseline > $OUT/ncu_launch_$TAG.log 2>&1; echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'texture_(forward_kernel|backward_stream_kernel)|unpack_grads' -s 6 -c 3 \
  -o $OUT/raster_$TAG -f python bench.py --steps 1 --warmup 1 --views-per-gpu 2 --no-ref-cuda --no-cpu-baseline > $OUT/ncu_full_$TAG.log 2>&1; echo "ncu full rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'tile_(count|colscan|scan|emit|local_sort)' -s 6 -c 6 \
  -o $OUT/sort_$TAG -f python bench.py --steps 1 --warmup 1 --views-per-gpu 2 --no-ref-cuda --no-cpu-baseline > $OUT/ncu_sort_$TAG.log 2>&1; echo "ncu sort rc=$?"
ls -la $OUT
