#!/bin/bash
OUT=gpurun_out/r2t
mkdir -p $OUT
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/tp_check.py > $OUT/tp_check.log 2>&1; echo "tp_check exit $?" >> $OUT/tp_check.log; tail -5 $OUT/tp_check.log
bash tools/r2_s.sh
