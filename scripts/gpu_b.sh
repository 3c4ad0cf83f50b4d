#!/bin/bash
# kernel iteration pass: targeted tests first (under a short timeout: a hung kernel must not hold the box), then
# the whole suite and a bench line.  Usage: bash scripts/gpu_b.sh <tag> "<pytest -k expression>"
TAG=${1:-r2_b}; KEXPR=${2:-dense}
OUT=gpurun_out; mkdir -p $OUT
timeout 300 python -m pytest tests/test_gpu_raster.py -x -q -k "$KEXPR" > $OUT/quick_$TAG.log 2>&1; rc=$?
echo "quick rc=$rc"; tail -25 $OUT/quick_$TAG.log
if [ $rc -ne 0 ]; then exit 0; fi
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?"
tail -12 $OUT/pytest_$TAG.log
timeout 600 python bench.py --steps 10 --warmup 3 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("$OUT/bench_$TAG.json"))
print({k:d[k] for k in ("value","ms_per_step","step_ms")}, d["e2e"]["value"], d["e2e_upstream_grads"]["value"])
print(d["roofline"]["per_entry_ms_per_view"]); print(d.get("parity")); print(d.get("reference_cuda_same_gpu"))
PY
tail -3 $OUT/bench_$TAG.err
