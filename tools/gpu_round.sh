#!/bin/bash
# One GPU-box visit: parity tests, the bench lines, the ncu launch list and full captures of the dominant kernels.
# Usage (from the repo root, under gpurun):  bash tools/gpu_round.sh [tag]
TAG=${1:-r01}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > $OUT/pytest_gpu.log 2>&1
echo "pytest exit $?" >> $OUT/pytest_gpu.log
( time timeout 900 python bench.py ) > $OUT/bench.log 2> $OUT/bench.err
echo "bench exit $?" >> $OUT/bench.err
( time timeout 600 python bench.py --impl reference --steps 2 --warmup 1 ) > $OUT/bench_ref.log 2> $OUT/bench_ref.err
timeout 300 python tools/bench_qmm.py > $OUT/qmm_q4.json 2>&1
timeout 300 python tools/bench_qmm.py --bits 8 > $OUT/qmm_q8.json 2>&1
timeout 300 python tools/bench_qmv.py --M 1 --graph 40 --pdl 1 > $OUT/qmv_m1.json 2>&1
timeout 300 python tools/bench_step_parts.py > $OUT/step_parts.json 2>&1
CMD="python bench.py --steps 2 --warmup 3 --prompt 8 --no-cpu-baseline --no-kernel-timing --no-prefill"
timeout 600 $CMD > $OUT/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:pie:: -c 1500 \
    --csv --log-file $OUT/launches.csv $CMD > $OUT/ncu_list.log 2>&1
timeout 600 $CMD > $OUT/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:qmv_kernel -s 400 -c 6 \
    -o $OUT/prof_qmv $CMD > $OUT/ncu_full.log 2>&1
timeout 300 python tools/bench_qmm.py --iters 1 > /dev/null 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:qmm_tc_kernel -s 8 -c 2 \
    -o $OUT/prof_qmm python tools/bench_qmm.py --iters 1 > $OUT/ncu_qmm.log 2>&1
# BASELINE configs 3 and 4 + captures of the kernels that dominate them
( timeout 400 python bench.py --bits 8 --batch 32 --prompt 2048 --paged --steps 32 --warmup 4 --no-cpu-baseline --no-kernel-timing --no-prefill ) > $OUT/config3.log 2> $OUT/config3.err
( timeout 400 python bench.py --bits 4 --batch 32 --prompt 2048 --steps 32 --warmup 4 --no-cpu-baseline --no-kernel-timing --no-prefill ) > $OUT/config3_q4.log 2>> $OUT/config3.err
timeout 300 python tools/bench_prefill.py > $OUT/config4.json 2>&1
timeout 300 python tools/bench_qmv.py --M 32 --bits 8 --graph 20 --pdl 1 > $OUT/qmv_m32_q8.json 2>&1
timeout 300 python tools/bench_qmv.py --M 32 --bits 4 --graph 20 --pdl 1 > $OUT/qmv_m32_q4.json 2>&1
PIE_PARTS="full;rope,attn" timeout 300 python tools/bench_step_parts.py --bits 8 --batch 32 --ctx 2048 > $OUT/step_parts_bs32_q8.json 2>&1
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:fa_prefill -s 2 -c 1 \
    -o $OUT/prof_fa_prefill python tools/bench_prefill.py --layers 2 > $OUT/ncu_fa_prefill.log 2>&1
PIE_PARTS="full" timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k 'regex:fa_decode|qmd_kernel' -s 40 -c 4 -o $OUT/prof_bs32 python tools/bench_step_parts.py --bits 8 --batch 32 --ctx 2048 --layers 2 > $OUT/ncu_bs32.log 2>&1
ls -la $OUT
tail -3 $OUT/pytest_gpu.log; cat $OUT/bench.log; tail -3 $OUT/bench.err; cat $OUT/bench_ref.log; cat $OUT/qmm_q4.json; cat $OUT/config3.log $OUT/config3_q4.log $OUT/config4.json; tail -2 $OUT/config3.err
