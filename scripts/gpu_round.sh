#!/bin/bash
# One GPU-box pass: parity tests, bench line, ncu launch list, ncu --set full (+ atomic counters) of the compositing and
# binning kernels.  Every ncu run follows a plain run of the same command that exited 0.
# Usage (from the repo root, under gpurun): bash scripts/gpu_round.sh <tag>
TAG=${1:-r2}
OUT=gpurun_out
mkdir -p $OUT
ATOM=lts__t_sectors_op_red.sum,lts__t_sectors_op_atom.sum,l1tex__t_set_accesses_pipe_lsu_mem_global_op_red.sum,launch__sm_count
python -m pytest tests -m gpu -x -q > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_$TAG.log
tail -4 $OUT/pytest_$TAG.log
python bench.py --steps 20 --warmup 3 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_ref_$TAG.json 2>> $OUT/bench_$TAG.err; echo "ref rc=$?"
python scripts/bench_loss.py 2>/dev/null | tail -1 > $OUT/bench_loss_$TAG.json
SMALL="python bench.py --steps 1 --warmup 1 --views-per-gpu 2 --no-ref-cuda --no-cpu-baseline"
$SMALL > $OUT/plain_$TAG.log 2>&1 || { echo "plain run failed"; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file $OUT/launches_$TAG.csv $SMALL > $OUT/ncu_launch_$TAG.log 2>&1; echo "ncu launches rc=$?"
ncu --set full --metrics $ATOM --clock-control none --import-source on -k regex:'texture_(forward_kernel|backward_stream_kernel)|unpack_grads' -s 6 -c 3 \
  -o $OUT/raster_$TAG -f $SMALL > $OUT/ncu_full_$TAG.log 2>&1; echo "ncu full rc=$?"
ncu --set full --metrics $ATOM --clock-control none --import-source on -k regex:'tile_(count|colscan|scan|emit|local_sort)|pack_records|project_bounds' -s 7 -c 7 \
  -o $OUT/sort_$TAG -f $SMALL > $OUT/ncu_sort_$TAG.log 2>&1; echo "ncu sort rc=$?"
ls -la $OUT | tail -12
