#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_raster.py -x -q -k "packed or padded or prepared or fused" 2>&1 | tail -4
for P in forward early forward early; do
timeout 300 python bench.py --steps 30 --warmup 3 --no-ref-cuda --no-cpu-baseline --pack $P > gpurun_out/bench_pack_$P.json 2> gpurun_out/bench_pack_$P.err || tail -5 gpurun_out/bench_pack_$P.err
python -c "
import json
d=json.loads(open('gpurun_out/bench_pack_$P.json').read().strip().splitlines()[-1])
print('$P', 'value %.1f'%d['value'], d['step_ms'], 'e2e %.1f'%d['e2e']['value'], d['e2e']['step_ms']['median'])
print('   ', {k: round(v,4) for k,v in d['roofline']['per_entry_ms_per_view'].items()})"
done
