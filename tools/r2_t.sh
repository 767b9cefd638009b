python -m pytest tests/test_gpu_attn_fused.py tests/test_gpu_model.py tests/test_gpu_configs.py -x -q 2>&1 | tail -2
python tools/r2_attn_warm.py 2>&1 | tail -3
python tools/timeline_step.py --layers 8 --dump gpurun_out/tl_l0.npy > gpurun_out/tl_l0.json 2>&1
python bench.py --no-cpu-baseline --no-prefill --no-kernel-timing 2>/dev/null | cut -c1-130
