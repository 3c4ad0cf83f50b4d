#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python bench.py --steps 30 --warmup 3 --no-ref-cuda --no-cpu-baseline > gpurun_out/bench_suite.json 2> gpurun_out/bench_suite.err || tail -5 gpurun_out/bench_suite.err
python -c "
import json
d=json.loads(open('gpurun_out/bench_suite.json').read().strip().splitlines()[-1])
print('value %.1f'%d['value'], d['step_ms'], 'e2e %.1f'%d['e2e']['value'], d['e2e']['step_ms'])
print('   ', {k: round(v,4) for k,v in d['roofline']['per_entry_ms_per_view'].items()})"
