"""Quantised KV cache, CPU side: the oracle's restatement of QuantizedKVCache.update_and_fetch
(PIE/cache/kv_cache/quantized.py:38-108) and quantized_scaled_dot_product_attention (PIE/models/base.py:56-89) against
tests/golden/ref_kvq.npz -- outputs of the reference's own compiled quantize / _qmm_t / _qmm kernels
(tools/make_golden_kvq.py)."""
import os

import numpy as np
import pytest

from oracle import oracle as O

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_kvq.npz"))
CASES = [tuple(int(v) for v in c) for c in GOLD["cases"]]


def _case(ci):
    g = lambda n: GOLD[f"c{ci}.{n}"]
    keys = (g("kq"), O.from_bf16_bits(g("ks")), O.from_bf16_bits(g("kb")))
    vals = (g("vq"), O.from_bf16_bits(g("vs")), O.from_bf16_bits(g("vb")))
    return g, keys, vals


@pytest.mark.parametrize("ci", range(len(CASES)))
def test_cache_rows_are_mx_quantize_of_the_bf16_rows(ci):
    g, keys, vals = _case(ci)
    c = O.QuantizedKVCache(64, 8)
    T = g("k").shape[2]
    cut = max(1, T // 3)
    c.update_and_fetch(g("k")[:, :, :cut], g("v")[:, :, :cut])          # prompt
    for t in range(cut, T):                                             # then one token at a time
        qk, qv = c.update_and_fetch(g("k")[:, :, t:t + 1], g("v")[:, :, t:t + 1])
    if cut == T:
        qk, qv = tuple(a[:, :, :T] for a in c.keys), tuple(a[:, :, :T] for a in c.values)
    assert c.offset == T and c.keys[0].shape[2] % 256 == 0
    for got, want in zip(qk + qv, keys + vals):
        assert np.array_equal(got, want)


@pytest.mark.parametrize("ci", range(len(CASES)))
def test_attention_matches_the_reference_kernels(ci):
    g, keys, vals = _case(ci)
    D = g("q").shape[-1]
    out_t = O.quantized_scaled_dot_product_attention(g("q"), keys, vals, D ** -0.5, None, 64, 8, mode="T")
    assert np.array_equal(out_t, g("out_bf16"))                          # the x86 bf16 kernels, bit for bit
    out_a = O.quantized_scaled_dot_product_attention(g("q"), keys, vals, D ** -0.5, None, 64, 8, mode="A")
    np.testing.assert_allclose(out_a, g("out_f32"), rtol=2e-5, atol=2e-5)  # float kernels (summation order differs)
    # (the two reference precisions themselves differ by up to ~0.8 here: the x86 scalar bf16 kernels round the running
    # sum to bf16 after every term; SURVEY appendix A -- the GPU path follows the fp32-accumulate semantics)


def test_attention_with_mask_and_grouped_heads():
    """L > 1: the additive causal mask of create_causal_mask goes through the same path (base.py:78-79)."""
    rng = np.random.default_rng(3)
    B, Hq, Hkv, T, D, L = 1, 4, 2, 9, 64, 4
    q = O.round_bf16(rng.standard_normal((B, Hq, L, D)).astype(np.float32))
    k = O.round_bf16(rng.standard_normal((B, Hkv, T, D)).astype(np.float32))
    v = O.round_bf16(rng.standard_normal((B, Hkv, T, D)).astype(np.float32))
    keys, vals = O.quantize(k, 64, 8), O.quantize(v, 64, 8)
    mask = O.create_causal_mask(L, T - L)
    got = O.quantized_scaled_dot_product_attention(q, keys, vals, D ** -0.5, mask, 64, 8, mode="A")
    kd, vd = O.dequantize(*keys, 64, 8, t_bf16=False), O.dequantize(*vals, 64, 8, t_bf16=False)
    want = O.scaled_dot_product_attention(q, kd, vd, D ** -0.5, mask, -1, "A")
    np.testing.assert_allclose(got, want, rtol=1e-4, atol=1e-4)


def test_transposed_false_shape_errors():
    w = np.zeros((8, 128), np.float32)
    wq, s, b = O.quantize(w, 64, 8)
    with pytest.raises(ValueError):
        O.quantized_matmul(np.zeros((2, 7), np.float32), wq, s, b, transpose=False, bits=8)
    with pytest.raises(ValueError):
        O.quantized_matmul(np.zeros((2, 8), np.float32), wq, s[:, :1], b[:, :1], transpose=False, bits=8)
