#!/bin/bash
OUT=gpurun_out/r2i
mkdir -p $OUT
nproc > $OUT/nproc.txt
( time timeout 2400 python -m pytest tests -m gpu -x -q ) > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_gpu.log; tail -15 $OUT/pytest_gpu.log
( timeout 600 python bench.py --no-cpu-baseline --no-prefill ) > $OUT/bench.log 2> $OUT/bench.err; echo "bench exit $?" >> $OUT/bench.err; cat $OUT/bench.log; tail -3 $OUT/bench.err
