#!/bin/bash
OUT=gpurun_out/r2s
mkdir -p $OUT
for n in 2; do
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 128 --warmup 8 --no-cpu-baseline ) > $OUT/bench_tp$n.log 2> $OUT/bench_tp$n.err; echo "N=$n exit $?"; grep "^{" $OUT/bench_tp$n.log > $OUT/bench_tp$n.json; python -c "
import sys, json
d = json.loads(open('$OUT/bench_tp$n.json').read())
print(d['value'], d['ms_per_step'], d['roofline']['avg_launch_us']); print(d.get('tp_exchange')); print(d.get('llama_3_1_70b_q4'))"
tail -3 $OUT/bench_tp$n.err
done
