python -m pytest tests/test_gpu_ops.py tests/test_gpu_model.py -x -q 2>&1 | tail -2
python tools/bench_qmv.py --M 1 --graph 40 --pdl 1 2>&1 | tail -1
PIE_LIB=$PWD/proxy_inference_engine_b200/libpie_b200_prev.so python tools/bench_qmv.py --M 1 --graph 40 --pdl 1 2>&1 | tail -1
python bench.py --no-cpu-baseline --no-prefill --no-kernel-timing 2>/dev/null | cut -c1-130
PIE_LIB=$PWD/proxy_inference_engine_b200/libpie_b200_prev.so python bench.py --no-cpu-baseline --no-prefill --no-kernel-timing 2>/dev/null | cut -c1-130
