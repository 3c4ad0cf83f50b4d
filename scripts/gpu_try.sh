#!/bin/bash
# Quick check of a kernel change: one pytest file (or -k filter) + per-entry timings.  Usage: bash scripts/gpu_try.sh <pytest args...>
OUT=gpurun_out; mkdir -p $OUT
echo "pytest: $(timeout 900 python -m pytest "$@" -x -q -m gpu 2>&1 | tail -1)"
bash scripts/gpu_dbg.sh 0 0 2>&1 | tail -2
