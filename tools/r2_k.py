import numpy as np, torch, sys, os, json
sys.path.insert(0, 'tests'); sys.path.insert(0, '.')
from parity_util import n2t_bf16, n2t_words, rel_l2, t2n
from oracle import oracle as orc
from proxy_inference_engine_b200 import ops
res = []
for bits in (4, 8):
    for (M, N, K) in ((16, 1024, 4096), (16, 2048, 4096), (16, 4096, 4096), (9, 4096, 4096), (32, 4096, 4096), (16, 4096, 2048),
                      (16, 4096, 1024), (32, 2368, 4096), (32, 2384, 4096), (16, 4096, 14336), (32, 28672, 4096)):
        rng = np.random.default_rng(M + N + K)
        w = (rng.uniform(-1, 1, (N, K)) / np.sqrt(K)).astype(np.float32)
        wq, s, b = orc.quantize(w, 64, bits)
        x = orc.round_bf16(rng.standard_normal((M, K)).astype(np.float32))
        ref = orc.quantized_matmul(x, wq, s, b, True, 64, bits, mode="B")
        got = t2n(ops.quantized_matmul(n2t_bf16(x), n2t_words(wq), n2t_bf16(s), n2t_bf16(b), True, 64, bits))
        bad_rows = [m for m in range(M) if rel_l2(got[m], ref[m]) > 1e-2]
        err_cols = np.nonzero(np.abs(got - ref).max(0) > 0.05 * np.abs(ref).max())[0]
        res.append((bits, M, N, K, round(rel_l2(got, ref), 4), bad_rows[:6], len(bad_rows), (int(err_cols.min()), int(err_cols.max()), len(err_cols)) if len(err_cols) else None))
        print(res[-1], flush=True)
