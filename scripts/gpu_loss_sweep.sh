#!/bin/bash
# Loss-kernel tile variants: build HERE, time on the box.
#   bash scripts/gpu_loss_sweep.sh build | run
MODE=$1
VARIANTS="32_8 24_6 24_8 16_4 16_8"
if [ "$MODE" = build ]; then
  mkdir -p gstex_b200/lib/variants; rm -f gstex_b200/lib/variants/lib_loss_*.so
  for v in $VARIANTS; do
    IFS=_ read TW O2 <<< "$v"
    touch gstex_b200/csrc/loss.cu
    GSTEX_B200_NVCC_EXTRA="-DGSTEX_LOSS_TILE_W=$TW -DGSTEX_LOSS_OUT2=$O2" python -c "from gstex_b200 import build; build.build()" > /tmp/build_$v.log 2>&1 || { echo "build $v failed"; tail -5 /tmp/build_$v.log; continue; }
    cp gstex_b200/lib/libgstex_b200.so gstex_b200/lib/variants/lib_loss_$v.so
    echo -n "$v: "; cuobjdump --dump-resource-usage gstex_b200/lib/libgstex_b200.so 2>/dev/null | grep -A1 "loss_.*ward_kernel" | grep -o "REG:[0-9]*" | tr '\n' ' '; echo
  done
  touch gstex_b200/csrc/loss.cu; python -c "from gstex_b200 import build; build.build()" > /dev/null 2>&1
else
  cp gstex_b200/lib/libgstex_b200.so /tmp/lib_default.so
  for v in $VARIANTS $VARIANTS; do
    cp gstex_b200/lib/variants/lib_loss_$v.so gstex_b200/lib/libgstex_b200.so
    echo -n "tileW_out2=$v  "; timeout 120 python scripts/bench_loss.py 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['per_entry_ms'], 'autograd ms', round(d['ms'],4))"
  done
  cp /tmp/lib_default.so gstex_b200/lib/libgstex_b200.so
fi
