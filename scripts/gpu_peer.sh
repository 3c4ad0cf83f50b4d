#!/bin/bash
# peer-memory all-reduce: tests at the box's GPU count, then bench A/B (peer / multimem / nccl) at N = all GPUs
# usage: bash scripts/gpu_peer.sh <N> <tag>
N=${1:-2}; TAG=${2:-peer}
mkdir -p gpurun_out
nvidia-smi topo -m 2>&1 | head -14
timeout 900 python -m pytest tests/test_gpu_peer_allreduce.py -x -q 2>&1 | tail -8
for AR in peer nccl multimem peer nccl; do
  P=$((29500 + RANDOM % 2000))
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P \
     bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline --allreduce $AR > gpurun_out/bench_${TAG}_n${N}_${AR}.json 2> gpurun_out/bench_${TAG}_n${N}_${AR}.err
  echo "rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_${TAG}_n${N}_${AR}.json').read().strip().splitlines()[-1])
    print('$AR', 'value %.1f'%d['value'], 'step', d['step_ms'], 'exchange', d.get('exchange_ms'), 'e2e %.1f'%d['e2e']['value'])
    print('   ', d['config']['parallelism'])
except Exception as e:
    print('$AR failed', e); print(open('gpurun_out/bench_${TAG}_n${N}_${AR}.err').read()[-1500:])
PY
done
