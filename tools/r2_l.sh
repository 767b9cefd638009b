#!/bin/bash
OUT=gpurun_out/r2l
mkdir -p $OUT
( time timeout 2400 python -m pytest tests -m gpu -q -s ) > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_gpu.log; grep -E "config1 parity|passed|failed|FAILED|real" $OUT/pytest_gpu.log | head -20
