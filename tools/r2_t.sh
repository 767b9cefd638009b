python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --no-cpu-baseline --no-prefill > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; tail -2 gpurun_out/bench_quick.err
python tools/timeline_step.py --layers 8 --dump gpurun_out/tl_j0.npy > gpurun_out/tl_j0.json 2>&1
