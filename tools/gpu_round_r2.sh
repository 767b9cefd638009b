#!/bin/bash
# GPU-box visits of round 2 (one profiler per visit):
#   bash tools/gpu_round_r2.sh [tag]        parity tests, bench lines (both arms), per-matrix GEMV table, step timeline, ncu launch list
#   bash tools/gpu_round_r2.sh [tag] full   `ncu --set full` capture of the decode GEMVs only
TAG=${1:-r02}
MODE=${2:-main}
OUT=gpurun_out/$TAG
mkdir -p $OUT
CMD="python bench.py --steps 2 --warmup 3 --prompt 8 --no-cpu-baseline --no-kernel-timing --no-prefill"
if [ "$MODE" = "full" ]; then
  timeout 600 $CMD > $OUT/plain2.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:qmv -s 500 -c 6 \
      -o $OUT/prof_qmv $CMD > $OUT/ncu_full.log 2>&1
  ls -la $OUT; tail -3 $OUT/ncu_full.log
  exit 0
fi
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
( time timeout 1800 python -m pytest tests -m gpu -q ) > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_gpu.log
( time timeout 900 python bench.py ) > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?" >> $OUT/bench.err
( timeout 600 python bench.py --impl reference --steps 2 --warmup 1 ) > $OUT/bench_reference.json 2> $OUT/bench_reference.err
timeout 300 python tools/bench_qmv.py --M 1 --graph 40 --pdl 1 > $OUT/qmv_m1.json 2>&1
PIE_QMV2=0 timeout 300 python tools/bench_qmv.py --M 1 --graph 40 --pdl 1 > $OUT/qmv_m1_cpasync_only.json 2>&1
timeout 300 python tools/bench_step_parts.py > $OUT/step_parts.json 2>&1
( timeout 400 python bench.py --bits 8 --batch 32 --prompt 2048 --paged --steps 32 --warmup 4 --no-cpu-baseline --no-kernel-timing --no-prefill ) > $OUT/config3_q8_bs32_paged.json 2> $OUT/config3.err
timeout 300 python tools/bench_prefill.py > $OUT/config4.json 2>&1
timeout 300 python tools/bench_attn_q8.py --splits 0 --bf16-splits 0 > $OUT/attn_bs32.json 2>&1
timeout 300 python tools/bench_step_parts.py --bits 8 --batch 32 --ctx 2048 > $OUT/step_parts_bs32.json 2>&1
timeout 300 python tools/bench_qmv.py --M 32 --bits 8 --graph 40 --pdl 1 --shapes qkv,o,gate_up,down > $OUT/qmi_m32.json 2>&1
timeout 300 python tools/timeline_step.py --layers 32 > $OUT/timeline_8b_ctx300.json 2> $OUT/timeline.err
timeout 600 $CMD > $OUT/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:pie:: -c 1500 \
    --csv --log-file $OUT/launches.csv $CMD > $OUT/ncu_list.log 2>&1
ls -la $OUT
tail -3 $OUT/pytest_gpu.log; cut -c1-400 $OUT/bench.json; tail -2 $OUT/bench.err; cut -c1-300 $OUT/bench_reference.json; cat $OUT/qmv_m1.json; cut -c1-200 $OUT/config3_q8_bs32_paged.json; cat $OUT/config4.json | tail -1
