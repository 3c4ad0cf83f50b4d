#!/bin/bash
# final-code check of the multi-rank path on 2 GPUs: tests, then bench.py the way the driver launches it
mkdir -p gpurun_out
true
for AR in auto multimem; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) \
    bench.py --gpus 2 --steps 30 --warmup 3 --allreduce $AR > gpurun_out/bench_n2_$AR.json 2> gpurun_out/bench_n2_$AR.err; echo "rc=$?"
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_n2_$AR.json').read().strip().splitlines()[-1])
    print('$AR value %.1f'%d['value'], d['step_ms'], 'exch', d.get('exchange_ms'), 'e2e %.1f'%d['e2e']['value'])
    print('   ', d['config']['parallelism']); print('   cpu', d.get('cpu_baseline'))
except Exception as e:
    print('failed', e); print(open('gpurun_out/bench_n2_$AR.err').read()[-2000:])
PY
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29640 \
  bench.py --impl reference --gpus 2 --steps 2 --warmup 1 | cut -c1-260
