#!/bin/bash
# Scaling pass on ONE 8-GPU box: the exchange microbenchmark with the multimem sweep, then N = 1, 2, 4, 8 back to back
# the way the driver launches them, then A/B of the exchange variants at N = 4 and 8.
TAG=${1:-r2}; MAXN=${2:-8}
OUT=gpurun_out; mkdir -p $OUT
summ() { python - "$1" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    print(sys.argv[1].split('/')[-1], "N", d["n_gpus"], "value", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 3), d["step_ms"],
          "| exch", (d.get("exchange_ms") or {}).get("median"), "| e2e", round(d["e2e"]["value"], 1), round(d["e2e"]["ms_per_step"], 3))
except Exception as e:
    print(sys.argv[1], "failed:", e)
PY
}
for N in 4 $MAXN; do
PEER_SWEEP=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2960$N scripts/peer_allreduce_bench.py 2>&1 | grep -v "^W\|^\*\*\*\|Warn\|OMP_NUM\|^$" > $OUT/peer_allreduce_${TAG}_n$N.txt
cat $OUT/peer_allreduce_${TAG}_n$N.txt | cut -c1-140
done
python bench.py --gpus 1 --steps 30 --warmup 3 --no-ref-cuda > $OUT/bench_${TAG}_n1.json 2> $OUT/bench_${TAG}_n1.err; summ $OUT/bench_${TAG}_n1.json
for N in 2 4 8; do
  [ $N -gt $MAXN ] && break
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N \
    bench.py --gpus $N --steps 30 --warmup 3 > $OUT/bench_${TAG}_n$N.json 2> $OUT/bench_${TAG}_n$N.err; summ $OUT/bench_${TAG}_n$N.json
done
for N in 4 $MAXN; do
for AR in nccl peer multimem; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$N \
    bench.py --gpus $N --steps 30 --warmup 3 --allreduce $AR --no-cpu-baseline > $OUT/bench_${TAG}_n${N}_$AR.json 2> $OUT/bench_${TAG}_n${N}_$AR.err
  summ $OUT/bench_${TAG}_n${N}_$AR.json
done
done
BENCH_PEER_CTAS=32 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $MAXN --master-addr 127.0.0.1 --master-port 29541 \
    bench.py --gpus $MAXN --steps 30 --warmup 3 --allreduce multimem --no-cpu-baseline > $OUT/bench_${TAG}_n${MAXN}_mm32.json 2> $OUT/bench_${TAG}_n${MAXN}_mm32.err
summ $OUT/bench_${TAG}_n${MAXN}_mm32.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $MAXN --master-addr 127.0.0.1 --master-port 29540 \
  bench.py --impl reference --gpus $MAXN --steps 2 --warmup 1 > $OUT/bench_ref_${TAG}_n$MAXN.json 2>> $OUT/bench_${TAG}_n$MAXN.err
cat $OUT/bench_ref_${TAG}_n$MAXN.json | cut -c1-300
nproc
