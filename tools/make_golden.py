"""Generate tests/golden/*.npz from the REFERENCE's own code (oracle/_ref/libpie_ref.so = MLX's
mlx/backend/cpu/quantized.cpp compiled where it lies under /root/reference, see oracle/Makefile).

Run in the build container (the reference tree is not on the GPU box):   python tools/make_golden.py
The fixtures pin (a) quantize<T,U> bytes (cpu/quantized.cpp:603-661) and (b) _qmm_t<float,bits,gs> outputs
(:96-153) on seeded inputs, and (c) the arange(512) case of MLX/tests/ops_tests.cpp:2944-2961.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle as orc  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def main():
    R = orc.ref()
    if R is None:
        raise SystemExit("oracle/_ref/libpie_ref.so missing: needs /root/reference mounted")
    os.makedirs(OUT, exist_ok=True)
    rng = np.random.default_rng(20261104)
    out = {}
    cases = []
    for bits in (2, 4, 8):
        for gs in (32, 64, 128):
            rows, K = 12, 512
            w = (rng.standard_normal((rows, K)) * rng.choice([1e-3, 0.05, 1.0, 30.0], (rows, 1))).astype(np.float32)
            w[3, :gs] = 0.0          # an all-zero group (round-trips exactly, test_quantized.py:24-28)
            w[4, gs:2 * gs] = 0.25   # a constant group (max == min)
            w[5, :gs] = np.abs(w[5, :gs])  # positive-only group
            w[6, :gs] = -np.abs(w[6, :gs])  # negative-only group
            tag = f"b{bits}g{gs}"
            # float32 quantize
            wq = np.empty((rows, K * bits // 32), np.uint32)
            s = np.empty((rows, K // gs), np.float32)
            b = np.empty_like(s)
            R.pie_ref_quantize_f32(w.reshape(-1), wq.reshape(-1), s.reshape(-1), b.reshape(-1), bits, gs, w.size)
            out[f"{tag}.w"], out[f"{tag}.f32.wq"], out[f"{tag}.f32.s"], out[f"{tag}.f32.b"] = w, wq, s, b
            # bfloat16 quantize (the checkpoint dtype)
            wb = orc.to_bf16_bits(w)
            wqb = np.empty_like(wq)
            sb = np.empty((rows, K // gs), np.uint16)
            bb = np.empty_like(sb)
            R.pie_ref_quantize_bf16(wb.reshape(-1), wqb.reshape(-1), sb.reshape(-1), bb.reshape(-1), bits, gs, w.size)
            out[f"{tag}.bf16.wq"], out[f"{tag}.bf16.s"], out[f"{tag}.bf16.b"] = wqb, sb, bb
            # _qmm_t<float> on the bf16-quantized weights, bf16-valued x (what the GPU path is fed)
            for M in (1, 3, 8):
                x = orc.round_bf16(rng.standard_normal((M, K)).astype(np.float32))
                y = np.empty((M, rows), np.float32)
                R.pie_ref_qmm_t_f32(y.reshape(-1), x.reshape(-1), wqb.reshape(-1),
                                    orc.from_bf16_bits(sb).reshape(-1), orc.from_bf16_bits(bb).reshape(-1),
                                    M, rows, K, bits, gs)
                out[f"{tag}.M{M}.x"], out[f"{tag}.M{M}.y"] = x, y
            cases.append(tag)
    # ops_tests.cpp:2944-2961: arange(512) x 4 rows at gs128
    ar = np.tile(np.arange(512, dtype=np.float32), (4, 1))
    for bits in (2, 4, 8):
        wq = np.empty((4, 512 * bits // 32), np.uint32)
        s = np.empty((4, 4), np.float32)
        b = np.empty_like(s)
        R.pie_ref_quantize_f32(ar.reshape(-1), wq.reshape(-1), s.reshape(-1), b.reshape(-1), bits, 128, ar.size)
        out[f"arange.b{bits}.wq"], out[f"arange.b{bits}.s"], out[f"arange.b{bits}.b"] = wq, s, b
    out["cases"] = np.array(cases)
    out["simd_max_size_f32"] = np.array([R.pie_ref_simd_max_size_f32()])
    path = os.path.join(OUT, "ref_quantized.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()
