# per-layer cost of decode attention: fused latency launch vs rope_kv_append + tensor-core decode attention
timeout 300 python -m pytest tests/test_gpu_attn_fused.py tests/test_gpu_ops.py -x -q -k "attn or attention or sdpa or paged" 2>&1 | tail -2
for cfg in ${CFGS:-"1 500" "1 2048" "1 8000" "2 2048" "4 500" "4 2048" "32 2048"}; do set -- $cfg
PIE_PARTS="full;rope,attn" timeout 150 python tools/bench_step_parts.py --bits 4 --batch $1 --ctx $2 --layers 8 --fused-attn 2>&1 | tail -1
PIE_PARTS="full;rope,attn" timeout 150 python tools/bench_step_parts.py --bits 4 --batch $1 --ctx $2 --layers 8 --unfused-attn 2>&1 | tail -1
done
