#!/bin/bash
# padded-texture (128-bit texel loads) check: tests, forward timing A/B, bench A/B
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_raster.py -x -q -k "padded or dense_forward" 2>&1 | tail -5
timeout 300 python scripts/fwd_time.py cfg5 0 2>&1 | tail -8
timeout 300 python scripts/fwd_time.py cfg4 0 2>&1 | tail -8
timeout 300 python bench.py --steps 10 --warmup 3 --no-ref-cuda --no-cpu-baseline > gpurun_out/bench_scalar.json 2> gpurun_out/bench_scalar.err
timeout 300 python bench.py --steps 10 --warmup 3 --no-ref-cuda --no-cpu-baseline --texel-loads 128 > gpurun_out/bench_128.json 2> gpurun_out/bench_128.err
timeout 300 python bench.py --steps 10 --warmup 3 --no-ref-cuda --no-cpu-baseline > gpurun_out/bench_scalar2.json 2>> gpurun_out/bench_scalar.err
timeout 300 python bench.py --steps 10 --warmup 3 --no-ref-cuda --no-cpu-baseline --texel-loads 128 > gpurun_out/bench_128b.json 2>> gpurun_out/bench_128.err
for f in scalar 128 scalar2 128b; do python -c "
import json,sys
d=json.loads(open('gpurun_out/bench_$f.json').read().strip().splitlines()[-1])
print('$f', d['value'], d['ms_per_step'], d['e2e']['value'])"; done
