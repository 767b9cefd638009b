#!/bin/bash
OUT=gpurun_out/r2o
mkdir -p $OUT
( timeout 900 python -m pytest tests/test_gpu_model.py -x -q -k "fused or parity" ) > $OUT/pytest.log 2>&1; echo "pytest exit $?" >> $OUT/pytest.log; tail -4 $OUT/pytest.log
for pf in 1 2; do PIE_STEP_PF=$pf timeout 300 python tools/trace_step.py > $OUT/trace_pf$pf.json 2> $OUT/trace.err; echo "pf=$pf"; cat $OUT/trace_pf$pf.json; tail -2 $OUT/trace.err; done
