#!/bin/bash
mkdir -p gpurun_out/r2k
for qmd in 0 1; do echo "PIE_QMD=$qmd"; PIE_QMD=$qmd timeout 900 python tools/r2_k.py 2>&1 | tee gpurun_out/r2k/qmd$qmd.log | tail -25; done
