#!/bin/bash
# Round-2 first GPU visit: TMA swizzle probe, parity of the TMA GEMV (bf16 and fp16 staging), micro-benchmarks.
OUT=gpurun_out/r2a
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
timeout 60 tools/probe/tma_probe > $OUT/tma_probe.txt 2>&1; echo "probe exit $?" >> $OUT/tma_probe.txt
head -40 $OUT/tma_probe.txt
for f16 in 0 1; do
  ( PIE_QMV2_F16=$f16 timeout 900 python -m pytest tests/test_gpu_ops.py tests/test_gpu_attn_fused.py -x -q -k "quantized_matmul or fused_norm" ) > $OUT/pytest_f16_$f16.log 2>&1
  echo "pytest f16=$f16 exit $?" >> $OUT/pytest_f16_$f16.log; tail -15 $OUT/pytest_f16_$f16.log
done
PIE_QMV2=0 timeout 300 python tools/bench_qmv.py --M 1 --graph 40 --pdl 1 > $OUT/qmv_old.json 2>&1; cat $OUT/qmv_old.json
PIE_QMV2_F16=0 timeout 300 python tools/bench_qmv.py --M 1 --graph 40 --pdl 1 > $OUT/qmv_new_bf16.json 2>&1; cat $OUT/qmv_new_bf16.json
PIE_QMV2_F16=1 timeout 300 python tools/bench_qmv.py --M 1 --graph 40 --pdl 1 > $OUT/qmv_new_f16.json 2>&1; cat $OUT/qmv_new_f16.json
( timeout 900 python -m pytest tests/test_gpu_model.py -x -q ) > $OUT/pytest_model.log 2>&1; echo "exit $?" >> $OUT/pytest_model.log; tail -5 $OUT/pytest_model.log
( timeout 600 python bench.py --no-cpu-baseline --no-prefill ) > $OUT/bench.log 2> $OUT/bench.err; echo "bench exit $?" >> $OUT/bench.err; cat $OUT/bench.log; tail -3 $OUT/bench.err
