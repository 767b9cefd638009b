"""Generate tests/golden/ref_kvq.npz: the quantised-KV attention of PIE/models/base.py:56-89 with every quantize /
quantized_matmul done by the REFERENCE's own compiled kernels (oracle/_ref/libpie_ref.so: quantize<T,U>, _qmm_t, _qmm of
MLX/mlx/backend/cpu/quantized.cpp, see oracle/ref_harness.cpp).  Only the softmax between the two matmuls is the
oracle's restatement (MLX's softmax is a graph primitive that cannot be built here).

Run in the build container (the reference tree is not on the GPU box):   python tools/make_golden_kvq.py
Per case: q [B, Hq, 1, D], k / v [B, Hkv, T, D] (bf16-valued), their quantised triples (bf16 arithmetic, 8 bits, groups
of 64), `out_bf16` (the reference's x86 bf16 kernels, every operation rounded) and `out_f32` (its float kernels on the
same quantised cache: the arithmetic a GPU path with fp32 accumulators should land on).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle as orc  # noqa: E402

CASES = [(1, 8, 2, 70, 128), (2, 4, 4, 33, 64), (1, 4, 1, 300, 128), (1, 2, 2, 1, 128)]


def ref_quantize_bf16(R, w, bits=8, gs=64):
    wb = orc.to_bf16_bits(w)
    wq = np.empty(w.shape[:-1] + (w.shape[-1] * bits // 32,), np.uint32)
    s = np.empty(w.shape[:-1] + (w.shape[-1] // gs,), np.uint16)
    b = np.empty_like(s)
    R.pie_ref_quantize_bf16(wb.reshape(-1), wq.reshape(-1), s.reshape(-1), b.reshape(-1), bits, gs, w.size)
    return wq, s, b


def main():
    R = orc.ref()
    if R is None:
        raise SystemExit("oracle/_ref/libpie_ref.so missing: needs /root/reference mounted")
    rng = np.random.default_rng(20261105)
    out = {}
    for ci, (B, Hq, Hkv, T, D) in enumerate(CASES):
        n_rep, bits, gs = Hq // Hkv, 8, 64
        scale = np.float32(D ** -0.5)
        q = orc.round_bf16(rng.standard_normal((B, Hq, 1, D)).astype(np.float32) * 2.0)
        k = orc.round_bf16((rng.standard_normal((B, Hkv, T, D)) * rng.choice([0.3, 1.0, 4.0], (B, Hkv, T, 1))).astype(np.float32))
        v = orc.round_bf16((rng.standard_normal((B, Hkv, T, D)) * rng.choice([0.1, 1.0, 3.0], (B, Hkv, 1, D))).astype(np.float32))
        kq, ks, kb = ref_quantize_bf16(R, k)
        vq, vs, vb = ref_quantize_bf16(R, v)
        o16 = np.empty((B, Hq, 1, D), np.float32)
        o32 = np.empty((B, Hq, 1, D), np.float32)
        for b_ in range(B):
            for h in range(Hkv):
                qs = q[b_, h * n_rep:(h + 1) * n_rep, 0] * scale                      # [n_rep, D]
                wk, sk, bk = kq[b_, h].reshape(-1), ks[b_, h].reshape(-1), kb[b_, h].reshape(-1)
                wv, sv, bv = vq[b_, h].reshape(-1), vs[b_, h].reshape(-1), vb[b_, h].reshape(-1)
                # bf16: queries * scale rounded; _qmm_t<bf16>; softmax (bf16 out); _qmm<bf16>
                x16 = orc.to_bf16_bits(qs)
                sc16 = np.empty((n_rep, T), np.uint16)
                R.pie_ref_qmm_t_bf16(sc16.reshape(-1), x16.reshape(-1), wk, sk, bk, n_rep, T, D, bits, gs)
                p16 = orc.to_bf16_bits(orc.softmax(orc.from_bf16_bits(sc16), "B"))
                y16 = np.empty((n_rep, D), np.uint16)
                R.pie_ref_qmm_n_bf16(y16.reshape(-1), p16.reshape(-1), wv, sv, bv, n_rep, D, T, bits, gs)
                o16[b_, h * n_rep:(h + 1) * n_rep, 0] = orc.from_bf16_bits(y16)
                # float kernels on the same quantised cache
                f = orc.from_bf16_bits
                sc = np.empty((n_rep, T), np.float32)
                R.pie_ref_qmm_t_f32(sc.reshape(-1), np.ascontiguousarray(qs, np.float32).reshape(-1), wk, f(sk), f(bk),
                                    n_rep, T, D, bits, gs)
                p = orc.softmax(sc, "A")
                y = np.empty((n_rep, D), np.float32)
                R.pie_ref_qmm_n_f32(y.reshape(-1), np.ascontiguousarray(p).reshape(-1), wv, f(sv), f(bv), n_rep, D, T,
                                    bits, gs)
                o32[b_, h * n_rep:(h + 1) * n_rep, 0] = y
        for name, arr in (("q", q), ("k", k), ("v", v), ("kq", kq), ("ks", ks), ("kb", kb), ("vq", vq), ("vs", vs),
                          ("vb", vb), ("out_bf16", o16), ("out_f32", o32)):
            out[f"c{ci}.{name}"] = arr
    out["cases"] = np.array(CASES)
    path = os.path.join(ROOT, "tests", "golden", "ref_kvq.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()
