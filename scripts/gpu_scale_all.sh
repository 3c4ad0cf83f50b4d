#!/bin/bash
# Scaling pass on ONE multi-GPU box, launched the way the driver does: N = 1, 2, 4, 8 back to back, then the
# binning-stream A/B at the largest N.  Usage (under gpurun --gpus 8): bash scripts/gpu_scale_all.sh <tag> [maxN]
TAG=${1:-r2}; MAXN=${2:-8}
OUT=gpurun_out; mkdir -p $OUT
summ() { python - "$1" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    print(sys.argv[1].split('/')[-1], "N", d["n_gpus"], "value", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 3), d["step_ms"],
          "| e2e", round(d["e2e"]["value"], 1), round(d["e2e"]["ms_per_step"], 3), "| e2e_up", round(d["e2e_upstream_grads"]["value"], 1),
          "| cpu", (d.get("cpu_baseline") or {}).get("cores"), round((d.get("cpu_baseline") or {}).get("value", 0), 3))
except Exception as e:
    print(sys.argv[1], "failed:", e)
PY
}
python bench.py --gpus 1 --steps 20 --warmup 3 --no-ref-cuda > $OUT/bench_${TAG}_n1.json 2> $OUT/bench_${TAG}_n1.err; summ $OUT/bench_${TAG}_n1.json
for N in 2 4 8; do
  [ $N -gt $MAXN ] && break
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N \
    bench.py --gpus $N --steps 20 --warmup 3 > $OUT/bench_${TAG}_n$N.json 2> $OUT/bench_${TAG}_n$N.err; summ $OUT/bench_${TAG}_n$N.json
done
for MODE in own current; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $MAXN --master-addr 127.0.0.1 --master-port 29530 \
    bench.py --gpus $MAXN --steps 20 --warmup 3 --binning-stream $MODE --no-cpu-baseline > $OUT/bench_${TAG}_n${MAXN}_$MODE.json 2> $OUT/bench_${TAG}_n${MAXN}_$MODE.err
  summ $OUT/bench_${TAG}_n${MAXN}_$MODE.json
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $MAXN --master-addr 127.0.0.1 --master-port 29540 \
  bench.py --impl reference --gpus $MAXN --steps 2 --warmup 1 > $OUT/bench_ref_${TAG}_n$MAXN.json 2>> $OUT/bench_${TAG}_n$MAXN.err
cat $OUT/bench_ref_${TAG}_n$MAXN.json | cut -c1-400
nproc
