#!/bin/bash
mkdir -p gpurun_out
for M in smi100 off smi200 nvml smi1000 off smi100; do
BENCH_CLOCKS=$M BENCH_DUMP_STEPS=1 timeout 300 python bench.py --steps 150 --warmup 3 --no-ref-cuda --no-cpu-baseline > gpurun_out/bench_clk_$M.json 2> gpurun_out/bench_clk_$M.err
python - <<PY
import json,re
d=json.loads(open('gpurun_out/bench_clk_$M.json').read().strip().splitlines()[-1])
err=open('gpurun_out/bench_clk_$M.err').read()
per=[float(x) for x in re.findall(r'per-step ms: (.*)',err)[0].split()]
out=[round(x,1) for x in per if x>10.6]
print('$M','value %.1f'%d['value'],'median %.2f'%d['step_ms']['median'],'outliers>10.6:',len(out),out[:12],'clocks',d.get('clocks'))
PY
done
