#!/bin/bash
OUT=gpurun_out/r2r
mkdir -p $OUT
( timeout 1800 python -m pytest tests -m gpu -x -q ) > $OUT/pytest.log 2>&1; echo "pytest exit $?" >> $OUT/pytest.log; tail -3 $OUT/pytest.log
timeout 300 python tools/bench_qmv.py --M 1 --graph 40 --pdl 1 > $OUT/qmv.json 2>&1; cat $OUT/qmv.json
( timeout 600 python bench.py --no-cpu-baseline --no-prefill ) > $OUT/bench.log 2> $OUT/bench.err; echo "bench exit $?" >> $OUT/bench.err; cut -c1-330 $OUT/bench.log; tail -2 $OUT/bench.err
