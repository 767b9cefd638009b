#!/bin/bash
# Tensor-parallel scaling on one 8-GPU box: Llama-3.1-70B q4 (and 8B) decode at TP 1/2/4/8, bs1 and bs16.
# Usage (under gpurun --gpus 8):  bash tools/tp_scaling.sh [tag]
TAG=${1:-tp}
OUT=gpurun_out/$TAG
mkdir -p $OUT
run() {  # model gpus batch port
  local m=$1 n=$2 b=$3 port=$4
  if [ "$n" = 1 ]; then
    CUDA_VISIBLE_DEVICES=0 timeout 240 python bench.py --model $m --batch $b --steps 64 --warmup 8 --no-cpu-baseline --no-prefill \
        > $OUT/${m}_tp${n}_bs${b}.json 2> $OUT/${m}_tp${n}_bs${b}.err
  else
    timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port \
        bench.py --gpus $n --model $m --batch $b --steps 64 --warmup 8 --no-cpu-baseline \
        > $OUT/${m}_tp${n}_bs${b}.log 2>&1
    grep -E "^\{" $OUT/${m}_tp${n}_bs${b}.log > $OUT/${m}_tp${n}_bs${b}.json
  fi
  echo "$m tp$n bs$b: $(cut -c1-130 $OUT/${m}_tp${n}_bs${b}.json)"
}
run llama-3.1-70b 8 1 29601
run llama-3.1-70b 4 1 29602
run llama-3.1-70b 2 1 29603
run llama-3.1-70b 1 1 0
run llama-3.1-70b 8 16 29604
run llama-3.1-8b 8 1 29605
run llama-3.1-8b 4 1 29606
