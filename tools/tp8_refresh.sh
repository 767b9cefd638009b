#!/bin/bash
# 8-GPU refresh of the tensor-parallel scaling numbers (70B q4 bs1 / bs16, 8B q4 bs1).
OUT=gpurun_out/tp8r; mkdir -p $OUT
run() { local m=$1 n=$2 b=$3 port=$4
  if [ "$n" = 1 ]; then
    CUDA_VISIBLE_DEVICES=0 timeout 300 python bench.py --model $m --batch $b --steps 64 --warmup 8 --no-cpu-baseline --no-prefill > $OUT/${m}_tp${n}_bs${b}.log 2>&1
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port \
      bench.py --gpus $n --model $m --batch $b --steps 64 --warmup 8 --no-cpu-baseline > $OUT/${m}_tp${n}_bs${b}.log 2>&1
  fi
  grep -E "^\{" $OUT/${m}_tp${n}_bs${b}.log > $OUT/${m}_tp${n}_bs${b}.json
  echo "$m tp$n bs$b: $(cut -c1-130 $OUT/${m}_tp${n}_bs${b}.json)"; }
run llama-3.1-70b 8 1 29801
run llama-3.1-70b 8 16 29802
run llama-3.1-70b 1 1 0
run llama-3.1-8b 8 1 29803
run llama-3.1-8b 4 1 29804
