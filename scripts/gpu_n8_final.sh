#!/bin/bash
# final-code run at N GPUs the way the driver launches it (one JSON line), for profiles/
N=${1:-8}; TAG=${2:-r2_final}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) \
    bench.py --gpus $N --steps 30 --warmup 3 > gpurun_out/bench_${TAG}_n$N.json 2> gpurun_out/bench_${TAG}_n$N.err; echo "rc=$?"
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_${TAG}_n$N.json').read().strip().splitlines()[-1])
    print('N', d['n_gpus'], 'value %.1f'%d['value'], d['step_ms'], 'exch', d.get('exchange_ms'), 'e2e %.1f'%d['e2e']['value'], d['e2e']['step_ms'])
    print('   ', d['config']['parallelism'])
except Exception as e:
    print('failed', e); print(open('gpurun_out/bench_${TAG}_n$N.err').read()[-2000:])
PY
