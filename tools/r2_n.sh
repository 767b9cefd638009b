#!/bin/bash
OUT=gpurun_out/r2n
mkdir -p $OUT
( time timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 128 --warmup 8 --no-cpu-baseline ) > $OUT/bench_tp2.log 2> $OUT/bench_tp2.err; echo "exit $?"; grep "^{" $OUT/bench_tp2.log | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print(d['value'], d['ms_per_step'], d['config']['parallelism']); print(d.get('tp_exchange')); print(d.get('llama_3_1_70b_q4'))"
tail -4 $OUT/bench_tp2.err
