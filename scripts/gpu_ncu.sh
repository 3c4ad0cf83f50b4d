#!/bin/bash
# ncu --set full (+ atomic counters) of selected kernels.  Usage: bash scripts/gpu_ncu.sh <tag> <kernel regex> [skip] [count]
TAG=$1; RE=$2; SKIP=${3:-4}; CNT=${4:-2}
OUT=gpurun_out; mkdir -p $OUT
ncu --set full --metrics lts__t_sectors_op_red.sum,lts__t_sectors_op_atom.sum,l1tex__t_set_accesses_pipe_lsu_mem_global_op_red.sum,lts__t_requests_srcunit_tex_op_red.sum,smsp__inst_executed_op_global_red.sum,launch__sm_count \
  --clock-control none --import-source on -k regex:"$RE" -s $SKIP -c $CNT \
  -o $OUT/ncu_$TAG -f python bench.py --steps 1 --warmup 1 --views-per-gpu 2 --no-ref-cuda --no-cpu-baseline > $OUT/ncu_$TAG.log 2>&1; echo "ncu rc=$?"
tail -3 $OUT/ncu_$TAG.log
