#!/bin/bash
# Build forward-kernel occupancy variants HERE (no GPU needed), then time them on the box:
#   bash scripts/gpu_fwd_sweep.sh build      (container)   -> gstex_b200/lib/variants/lib_W_C_B.so
#   bash scripts/gpu_fwd_sweep.sh run cfg5   (under gpurun)
MODE=$1; CFG=${2:-cfg5}
VARIANTS="5_4_32 5_4_16 5_5_16 6_4_16 4_6_16 8_3_16 7_3_32"
if [ "$MODE" = build ]; then
  mkdir -p gstex_b200/lib/variants; cp gstex_b200/lib/libgstex_b200.so /tmp/lib_default.so
  for v in $VARIANTS; do
    IFS=_ read W C B <<< "$v"
    touch gstex_b200/csrc/raster_fwd.cu
    GSTEX_B200_NVCC_EXTRA="-DGSTEX_FWD_WARPS=$W -DGSTEX_FWD_CTAS=$C -DGSTEX_FWD_BATCH=$B" python -m gstex_b200.build > /dev/null 2>&1 || echo "build $v failed"
    cp gstex_b200/lib/libgstex_b200.so gstex_b200/lib/variants/lib_$v.so
    echo -n "$v: "; cuobjdump --dump-resource-usage gstex_b200/lib/libgstex_b200.so 2>/dev/null | grep -A1 "texture_forward_kernelILi3ELb1ELb0" | grep -o "REG:[0-9]*\|STACK:[0-9]*" | tr '\n' ' '; echo
  done
  touch gstex_b200/csrc/raster_fwd.cu; python -m gstex_b200.build > /dev/null 2>&1
else
  cp gstex_b200/lib/libgstex_b200.so /tmp/lib_default.so
  for v in $VARIANTS; do
    cp gstex_b200/lib/variants/lib_$v.so gstex_b200/lib/libgstex_b200.so
    echo -n "W_C_B=$v  "; timeout 120 python scripts/fwd_time.py $CFG 0 2>&1 | grep "record 1" | sed 's/.*median/median/'
  done
  cp /tmp/lib_default.so gstex_b200/lib/libgstex_b200.so
fi
