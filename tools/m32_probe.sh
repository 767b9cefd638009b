# batch 9..32 decode: GEMM microbench per shape and the whole step (config 3 = q8 bs32 ctx2048)
timeout 120 python tools/bench_qmv.py --M 32 --bits 4 --graph 20 --pdl 1 2>&1 | tail -1
timeout 120 python tools/bench_qmv.py --M 16 --bits 4 --graph 20 --pdl 1 2>&1 | tail -1
timeout 120 python tools/bench_qmv.py --M 32 --bits 8 --graph 20 --pdl 1 2>&1 | tail -1
PIE_QMD=0 timeout 120 python tools/bench_qmv.py --M 32 --bits 8 --graph 20 --pdl 1 2>&1 | tail -1
PIE_PARTS="full;rope,attn" timeout 150 python tools/bench_step_parts.py --bits 8 --batch 32 --ctx 2048 2>&1 | tail -1
PIE_PARTS="full;rope,attn" timeout 150 python tools/bench_step_parts.py --bits 4 --batch 32 --ctx 2048 2>&1 | tail -1
PIE_PARTS="full" timeout 150 python tools/bench_step_parts.py --bits 4 --batch 16 --ctx 500 2>&1 | tail -1
