#!/bin/bash
mkdir -p gpurun_out
timeout 600 python bench.py --config cfg4 --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/bench_cfg4_b.json 2> gpurun_out/bench_cfg4_b.err; echo rc=$?
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_cfg4_b.json').read().strip().splitlines()[-1])
print('value %.1f'%d['value'], 'ms/view %.3f'%d['config']['ms_per_view'], d['step_ms'], 'e2e %.1f'%d['e2e']['value'])
print(d.get('reference_cuda_same_gpu')); print({k:v for k,v in (d.get('parity') or {}).items() if k!='what'})
print({k: round(v,3) for k,v in d['roofline']['per_entry_ms_per_view'].items()})
PY
