#!/bin/bash
# diagnosis of step outliers: per-step times + allocator traffic of every timed leg (BENCH_CLOCKS=off|smi<ms> optional)
mkdir -p gpurun_out
for M in a b c; do
BENCH_DUMP_STEPS=1 timeout 300 python bench.py --steps ${STEPS:-150} --warmup 3 --no-ref-cuda --no-cpu-baseline > gpurun_out/bench_clk_$M.json 2> gpurun_out/bench_clk_$M.err
python - <<PY
import json,re
d=json.loads(open('gpurun_out/bench_clk_$M.json').read().strip().splitlines()[-1])
err=open('gpurun_out/bench_clk_$M.err').read()
for i,l in enumerate(re.findall(r'per-step ms: (.*)',err)):
    per=[float(x) for x in l.split()]
    out=[(j,round(x,1)) for j,x in enumerate(per) if x>1.06*sorted(per)[len(per)//2]]
    print('$M leg',i,'median %.2f'%sorted(per)[len(per)//2],'outliers:',len(out),out[:12])
for l in re.findall(r'cudaMalloc.*',err): print('  ',l)
print('value %.1f'%d['value'],'e2e %.1f'%d['e2e']['value'])
PY
done
