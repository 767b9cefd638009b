# refresh of the ncu launch list and the prefill figure on the final code of the round
OUT=gpurun_out/final; mkdir -p $OUT
timeout 300 python tools/bench_prefill.py > $OUT/config4.json 2>&1
timeout 200 python tools/bench_step_parts.py > $OUT/step_parts.json 2>&1
CMD="python bench.py --steps 2 --warmup 3 --prompt 8 --no-cpu-baseline --no-kernel-timing --no-prefill"
timeout 600 $CMD > $OUT/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:pie:: -c 1500 \
    --csv --log-file $OUT/launches.csv $CMD > $OUT/ncu_list.log 2>&1
tail -1 $OUT/config4.json; tail -1 $OUT/step_parts.json; wc -l $OUT/launches.csv
