#!/bin/bash
# 70B q4 decode on 4 GPUs (tensor parallel 4 and 2): refresh of the scaling numbers after kernel changes.
OUT=gpurun_out/tp4r; mkdir -p $OUT
run() { local m=$1 n=$2 b=$3 port=$4
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port \
      bench.py --gpus $n --model $m --batch $b --steps 64 --warmup 8 --no-cpu-baseline > $OUT/${m}_tp${n}_bs${b}.log 2>&1
  grep -E "^\{" $OUT/${m}_tp${n}_bs${b}.log > $OUT/${m}_tp${n}_bs${b}.json
  echo "$m tp$n bs$b: $(cut -c1-130 $OUT/${m}_tp${n}_bs${b}.json)"; }
run llama-3.1-70b 4 1 29701
run llama-3.1-70b 4 16 29702
run llama-3.1-70b 2 1 29703
