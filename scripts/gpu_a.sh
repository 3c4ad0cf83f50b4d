#!/bin/bash
# round-2 first pass: full-size parity, whole GPU suite, bench lines
TAG=${1:-r2_a}
OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests/test_gpu_fullsize.py -q -s > $OUT/fullsize_$TAG.log 2>&1; echo "fullsize rc=$?"
tail -60 $OUT/fullsize_$TAG.log
python -m pytest tests -m gpu -q --deselect tests/test_gpu_fullsize.py > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?"
tail -15 $OUT/pytest_$TAG.log
python bench.py --steps 10 --warmup 3 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench rc=$?"
cat $OUT/bench_$TAG.json; tail -5 $OUT/bench_$TAG.err
