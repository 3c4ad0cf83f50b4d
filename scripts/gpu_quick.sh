#!/bin/bash
# Quick GPU pass: parity tests + one bench line.  Usage: bash scripts/gpu_quick.sh <tag> [extra bench args]
TAG=${1:-q}; shift
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests -m gpu -x -q > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?"
tail -15 $OUT/pytest_$TAG.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline "$@" > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench rc=$?"
tail -5 $OUT/bench_$TAG.err
python - <<PY
import json
d=json.load(open("$OUT/bench_$TAG.json"))
print("value",d["value"],"ms/view",d["config"]["ms_per_view"],"e2e",d["e2e"]["value"])
print(json.dumps(d["roofline"]["per_entry_ms_per_view"],indent=0))
print(d.get("reference_cuda_same_gpu"))
PY
