python -m pytest tests/test_gpu_ops.py tests/test_gpu_model.py -x -q 2>&1 | tail -2
python tools/timeline_step.py --layers 8 --dump gpurun_out/tl_m0.npy > gpurun_out/tl_m0.json 2>&1
python bench.py --no-cpu-baseline --no-prefill --no-kernel-timing 2>/dev/null | cut -c1-130
python bench.py --no-cpu-baseline --no-prefill --no-kernel-timing 2>/dev/null | cut -c1-130
