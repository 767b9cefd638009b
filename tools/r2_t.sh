python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python tools/bench_qmv.py --bits 8 --M 32 --graph 20 2>&1 | tail -1 | cut -c1-600
python bench.py --bits 8 --batch 32 --prompt 2048 --paged --steps 32 --warmup 4 --no-cpu-baseline --no-kernel-timing --no-prefill 2>/dev/null | cut -c1-140
