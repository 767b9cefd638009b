#!/bin/bash
OUT=gpurun_out/r2q
mkdir -p $OUT
( timeout 900 python -m pytest tests/test_gpu_ops.py tests/test_gpu_attn_fused.py tests/test_gpu_model.py -x -q -k "quantized_matmul or fused or parity" ) > $OUT/pytest.log 2>&1; echo "pytest exit $?" >> $OUT/pytest.log; tail -3 $OUT/pytest.log
PIE_QMV2=2 timeout 300 python tools/bench_qmv.py --M 1 --graph 40 --pdl 1 > $OUT/qmv_all.json 2>&1; cat $OUT/qmv_all.json
PIE_QMV2=2 PIE_QMV2_F16=0 timeout 300 python tools/bench_qmv.py --M 1 --graph 40 --pdl 1 > $OUT/qmv_all_bf16.json 2>&1; cat $OUT/qmv_all_bf16.json
PIE_STEP_PF=2 timeout 300 python tools/trace_step.py > $OUT/trace.json 2> $OUT/trace.err; cat $OUT/trace.json; tail -2 $OUT/trace.err
