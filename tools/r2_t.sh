python -m pytest tests/test_gpu_qmi.py -x -q 2>&1 | tail -2
for M in 1 4 8; do
python tools/bench_qmv.py --bits 8 --M $M --graph 20 --pdl 1 2>&1 | tail -1 | cut -c1-500
PIE_QMI=2 python tools/bench_qmv.py --bits 8 --M $M --graph 20 --pdl 1 2>&1 | tail -1 | cut -c1-500
done
