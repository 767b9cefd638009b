python -m pytest tests/test_zz_no_alloc.py -x -q 2>&1 | tail -3
python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python tools/timeline_step.py --layers 8 --dump gpurun_out/tl_h1.npy > gpurun_out/tl_h1.json 2>&1
PIE_QMV2_DBG=8 python tools/timeline_step.py --layers 8 --dump gpurun_out/tl_h1d8.npy > gpurun_out/tl_h1d8.json 2>&1
