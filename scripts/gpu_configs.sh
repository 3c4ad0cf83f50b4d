#!/bin/bash
# The other BASELINE.json configurations at full size (single GPU): our path + the reference CUDA kernels beside it.
# Usage: bash scripts/gpu_configs.sh <tag>
TAG=${1:-cfgs}; OUT=gpurun_out; mkdir -p $OUT
for c in ${CFGS:-cfg2 cfg2_2dgs cfg3 cfg4}; do
  V=4; [ $c = cfg4 ] && V=8
  timeout 900 python bench.py --config $c --steps 3 --warmup 3 --views-per-gpu $V --no-cpu-baseline > $OUT/bench_${TAG}_$c.json 2> $OUT/bench_${TAG}_$c.err; echo "$c rc=$?"
  python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_${TAG}_$c.json"))
    r=d.get("reference_cuda_same_gpu") or {}
    print("$c", "ms/view", round(d["config"]["ms_per_view"],3), "Mpix/s", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "I", d["config"]["intersections_view0"], "ref_cuda ms/view", r.get("ms_per_view"), r.get("error"))
    print({k: round(v,3) for k,v in d["roofline"]["per_entry_ms_per_view"].items()})
except Exception as e:
    print("$c failed", e); print(open("$OUT/bench_${TAG}_$c.err").read()[-1500:])
PY
done
