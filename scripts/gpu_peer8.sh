#!/bin/bash
# 8-GPU (or N) check of the exchange step: microbenchmark, then bench.py with our kernel and with NCCL
N=${1:-8}; TAG=${2:-peer}
mkdir -p gpurun_out
P=$((29500 + RANDOM % 2000))
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P scripts/peer_allreduce_bench.py 2>&1 | grep -v "^W\|^\*\*\*\|Warn" | tee gpurun_out/peer_allreduce_${TAG}_n${N}.txt | tail -12
for AR in ${3:-peer nccl multimem}; do
  P=$((29500 + RANDOM % 2000))
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P \
     bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline --allreduce $AR > gpurun_out/bench_${TAG}_n${N}_${AR}.json 2> gpurun_out/bench_${TAG}_n${N}_${AR}.err
  echo "rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_${TAG}_n${N}_${AR}.json').read().strip().splitlines()[-1])
    print('$AR', 'value %.1f'%d['value'], 'step', d['step_ms'], 'exchange', d.get('exchange_ms'), 'e2e %.1f'%d['e2e']['value'])
    pe=d['roofline']['per_entry_ms_per_view']
    print('   ', {k: round(v*8,4) for k,v in pe.items()})
except Exception as e:
    print('$AR failed', e); print(open('gpurun_out/bench_${TAG}_n${N}_${AR}.err').read()[-1500:])
PY
done
