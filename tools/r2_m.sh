#!/bin/bash
OUT=gpurun_out/r2m
mkdir -p $OUT
nvidia-smi -L > $OUT/gpus.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/tp_check.py > $OUT/tp_check.log 2>&1; echo "tp_check exit $?" >> $OUT/tp_check.log; tail -6 $OUT/tp_check.log
for os in 1; do
PIE_TP_ONESHOT=$os timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 128 --warmup 8 --no-cpu-baseline --no-prefill > $OUT/bench_tp2_oneshot$os.log 2> $OUT/bench_tp2_oneshot$os.err; echo "exit $?"; tail -2 $OUT/bench_tp2_oneshot$os.log | cut -c1-600; tail -3 $OUT/bench_tp2_oneshot$os.err
done
