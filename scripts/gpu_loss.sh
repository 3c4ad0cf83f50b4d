#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_loss.py -x -q 2>&1 | tail -5
timeout 300 python scripts/bench_loss.py 2>&1 | tail -3
timeout 300 python bench.py --steps 10 --warmup 3 --no-ref-cuda --no-cpu-baseline > gpurun_out/bench_loss3.json 2> gpurun_out/bench_loss3.err
python -c "
import json
d=json.loads(open('gpurun_out/bench_loss3.json').read().strip().splitlines()[-1])
print('value', d['value'], 'ms/step', d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'])"
