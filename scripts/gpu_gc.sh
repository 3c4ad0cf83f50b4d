#!/bin/bash
mkdir -p gpurun_out
for G in default freeze default freeze; do
BENCH_GC=$G BENCH_DUMP_STEPS=1 timeout 300 python bench.py --steps 60 --warmup 3 --no-ref-cuda --no-cpu-baseline > gpurun_out/bench_gc_$G.json 2> gpurun_out/bench_gc_$G.err
echo "== $G"; grep "per-step" gpurun_out/bench_gc_$G.err | head -2 | cut -c1-700
python -c "
import json
d=json.loads(open('gpurun_out/bench_gc_$G.json').read().strip().splitlines()[-1])
print('value', d['value'], d['step_ms'], 'e2e', d['e2e']['value'], d['e2e']['step_ms'])"
done
