"""Decode attention over the quantised KV cache (pie_decode_attn_q8) beside the bf16 kernels (pie_decode_attn), 8B
shapes, paged, per-layer microseconds and achieved GB/s of KV bytes.  Two cache sets are alternated so that every launch
reads more than the 126 MB L2 holds.   python tools/bench_attn_q8.py [--batch 32] [--ctx 2048] [--splits 0]"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from proxy_inference_engine_b200 import _lib, ops  # noqa: E402
from proxy_inference_engine_b200.cache import PagedKVCache, PagedQuantizedKVCache  # noqa: E402


def timed(fn, iters=20, warm=5):
    for _ in range(warm):
        fn(0), fn(1)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(iters):
        fn(i & 1)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--ctx", type=int, default=2048)
    ap.add_argument("--splits", type=int, nargs="*", default=[0])
    ap.add_argument("--bf16-splits", type=int, nargs="*", default=[0])
    ap.add_argument("--hq", type=int, default=32)
    ap.add_argument("--hkv", type=int, default=8)
    ap.add_argument("--dim", type=int, default=128)
    a = ap.parse_args()
    B, T, Hq, Hkv, D = a.batch, a.ctx, a.hq, a.hkv, a.dim
    n_pg = (T + 63) // 64
    g = torch.Generator().manual_seed(0)
    table = torch.randperm(B * n_pg, generator=g).reshape(B, n_pg).to(torch.int32).cuda()
    pos = torch.full((B,), T - 1, dtype=torch.int32, device="cuda")
    q = torch.randn((B, Hq, D), device="cuda").to(torch.bfloat16)
    sets_q, sets_b = [], []
    for _ in range(2):
        c = PagedQuantizedKVCache(B * n_pg, Hkv, D, table)
        for t in c.keys + c.values:
            if t.dtype == torch.int32:
                t.copy_(torch.randint(-2**31, 2**31 - 1, t.shape, device="cuda", dtype=torch.int64).to(torch.int32))
            else:
                t.copy_(torch.randn(t.shape, device="cuda") * 0.02)
        sets_q.append((c, c.desc()))
        kb = torch.randn((B * n_pg, 64, Hkv, D), device="cuda").to(torch.bfloat16)
        vb = torch.randn((B * n_pg, 64, Hkv, D), device="cuda").to(torch.bfloat16)
        pc = PagedKVCache(kb, vb, table)
        sets_b.append((pc, pc.layout()))
    out = torch.empty((B, Hq * D), dtype=torch.bfloat16, device="cuda")
    rec = {"batch": B, "ctx": T, "Hq": Hq, "Hkv": Hkv, "D": D, "rows": []}
    bytes_q = B * Hkv * T * 2 * (D + 4 * (D // 64))
    bytes_b = B * Hkv * T * 2 * D * 2
    L = _lib.load()
    for ns in a.splits:
        nq = ns or L.pie_decode_attn_q8_pick_splits(B, Hq, Hkv, T)
        us = timed(lambda i: ops.decode_attention_q8(q, sets_q[i][1], pos, D ** -0.5, Hkv, nsplit=nq, out=out))
        rec["rows"].append({"kernel": "q8", "nsplit": nq, "us": round(us, 2), "GBps": round(bytes_q / us / 1e3, 1)})
    for ns in a.bf16_splits:
        nb = ns or L.pie_decode_attn_pick_splits(B, Hq, Hkv, T)
        us = timed(lambda i: ops.decode_attention(q, sets_b[i][0].k_pool, sets_b[i][0].v_pool, sets_b[i][1], pos, D ** -0.5,
                                                  nsplit=nb, max_context=T))
        rec["rows"].append({"kernel": "bf16 (tensor-core flash-decode)", "nsplit": nb, "us": round(us, 2),
                            "GBps": round(bytes_b / us / 1e3, 1)})
    rec["kv_bytes"] = {"q8": bytes_q, "bf16": bytes_b}
    print(json.dumps(rec))


if __name__ == "__main__":
    main()
