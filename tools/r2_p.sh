#!/bin/bash
OUT=gpurun_out/r2p
mkdir -p $OUT
CMD="python tools/trace_step.py --layers 8"
timeout 300 $CMD > $OUT/plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:decode_step2 -s 2 -c 1 -o $OUT/prof_step2 $CMD > $OUT/ncu.log 2>&1
tail -2 $OUT/ncu.log
