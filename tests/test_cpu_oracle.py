"""CPU: pin the oracle (oracle/pie_oracle.c restatement) against the golden vectors generated from the reference's own
compiled code (tools/make_golden.py -> tests/golden/ref_quantized.npz) and against the property tests MLX holds for
this path (MLX/python/tests/test_quantized.py, test_fast.py, test_fast_sdpa.py)."""
import os

import numpy as np
import pytest

from oracle import oracle as orc

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_quantized.npz"))
CASES = [str(c) for c in GOLD["cases"]]


def _bg(tag):
    b, g = tag[1:].split("g")
    return int(b), int(g)


@pytest.mark.parametrize("tag", CASES)
def test_quantize_bit_exact_vs_reference_f32(tag):
    bits, gs = _bg(tag)
    wq, s, b = orc.quantize(GOLD[f"{tag}.w"], gs, bits, dtype="f32")
    assert np.array_equal(wq, GOLD[f"{tag}.f32.wq"])
    assert np.array_equal(s.view(np.uint32), GOLD[f"{tag}.f32.s"].view(np.uint32))
    assert np.array_equal(b.view(np.uint32), GOLD[f"{tag}.f32.b"].view(np.uint32))


@pytest.mark.parametrize("tag", CASES)
def test_quantize_bit_exact_vs_reference_bf16(tag):
    bits, gs = _bg(tag)
    wq, s, b = orc.quantize(GOLD[f"{tag}.w"], gs, bits, dtype="bf16")
    assert np.array_equal(wq, GOLD[f"{tag}.bf16.wq"])
    assert np.array_equal(orc.to_bf16_bits(s), GOLD[f"{tag}.bf16.s"])
    assert np.array_equal(orc.to_bf16_bits(b), GOLD[f"{tag}.bf16.b"])


@pytest.mark.parametrize("tag", CASES)
@pytest.mark.parametrize("M", [1, 3, 8])
def test_qmm_t_sequential_bit_exact_vs_reference(tag, M):
    """orc_qmm_t_seq keeps _qmm_t's strict k order (cpu/quantized.cpp:116-148): float32 results equal bit for bit."""
    bits, gs = _bg(tag)
    s, b = orc.from_bf16_bits(GOLD[f"{tag}.bf16.s"]), orc.from_bf16_bits(GOLD[f"{tag}.bf16.b"])
    y = orc.quantized_matmul(GOLD[f"{tag}.M{M}.x"], GOLD[f"{tag}.bf16.wq"], s, b, True, gs, bits, mode="A",
                             sequential=True)
    assert np.array_equal(y.view(np.uint32), GOLD[f"{tag}.M{M}.y"].view(np.uint32))


@pytest.mark.parametrize("tag", CASES)
def test_qmm_t_fast_within_reassociation(tag):
    """The blocked (reassociated) oracle used for whole-model runs stays within fp32 reassociation of _qmm_t."""
    bits, gs = _bg(tag)
    s, b = orc.from_bf16_bits(GOLD[f"{tag}.bf16.s"]), orc.from_bf16_bits(GOLD[f"{tag}.bf16.b"])
    y = orc.quantized_matmul(GOLD[f"{tag}.M8.x"], GOLD[f"{tag}.bf16.wq"], s, b, True, gs, bits, mode="A")
    ref = GOLD[f"{tag}.M8.y"]
    assert np.abs(y - ref).max() <= 2e-5 * max(1.0, np.abs(ref).max())


@pytest.mark.parametrize("bits", [2, 4, 8])
def test_arange_case_of_ops_tests(bits):
    """MLX/tests/ops_tests.cpp:2944-2961: shapes + |x - dq(x)| <= 127 / 2^bits; plus bytes equal to the reference."""
    ar = np.tile(np.arange(512, dtype=np.float32), (4, 1))
    wq, s, b = orc.quantize(ar, 128, bits, dtype="f32")
    assert wq.shape == (4, 512 * bits // 32) and s.shape == (4, 4) and b.shape == (4, 4)
    assert np.array_equal(wq, GOLD[f"arange.b{bits}.wq"])
    assert np.array_equal(s, GOLD[f"arange.b{bits}.s"]) and np.array_equal(b, GOLD[f"arange.b{bits}.b"])
    dq = orc.dequantize(wq, s, b, 128, bits, t_bf16=False)
    assert np.abs(dq - ar).max() <= 127.0 / (1 << bits)


def test_reference_takes_scalar_path_on_x86():
    # simd::max_size<float> == 1 off Apple -> _qmm_t (not _qmm_t_simd) is what MLX-CPU runs here
    assert int(GOLD["simd_max_size_f32"][0]) == 1


def test_live_reference_matches_golden_when_present():
    """In the build container the compiled reference is present: the committed fixtures must be reproducible."""
    R = orc.ref()
    if R is None:
        pytest.skip("oracle/_ref not built (no reference tree on this box)")
    tag = "b4g64"
    w = GOLD[f"{tag}.w"]
    wq = np.empty_like(GOLD[f"{tag}.f32.wq"]); s = np.empty_like(GOLD[f"{tag}.f32.s"]); b = np.empty_like(s)
    R.pie_ref_quantize_f32(w.reshape(-1), wq.reshape(-1), s.reshape(-1), b.reshape(-1), 4, 64, w.size)
    assert np.array_equal(wq, GOLD[f"{tag}.f32.wq"]) and np.array_equal(s, GOLD[f"{tag}.f32.s"])


# ---------------------------------------------------------------- MLX property tests, ported (test_quantized.py)
@pytest.mark.parametrize("gs", [32, 64, 128])
@pytest.mark.parametrize("bits", [2, 4, 8])
def test_quantize_dequantize_roundtrip(gs, bits):
    """test_quantized.py:11-28."""
    rng = np.random.default_rng(0)
    w = rng.standard_normal((128, 512)).astype(np.float32)
    wq, s, b = orc.quantize(w, gs, bits, dtype="f32")
    w_hat = orc.dequantize(wq, s, b, gs, bits, t_bf16=False)
    eps = 1e-6
    assert ((np.abs(w - w_hat).reshape(128, -1, gs)) <= (np.abs(s)[..., None] + eps)).all()
    w0 = np.zeros((128, 512), np.float32)
    wq, s, b = orc.quantize(w0, gs, bits, dtype="f32")
    assert (orc.dequantize(wq, s, b, gs, bits, t_bf16=False) == 0).all()


@pytest.mark.parametrize("gs,bits,M,N,K", [(64, 4, 8, 128, 256), (64, 8, 33, 256, 128), (32, 2, 32, 128, 128),
                                            (128, 4, 64, 256, 256), (64, 4, 1, 33, 128), (64, 8, 2, 3, 64)])
def test_qmm_vs_dense(gs, bits, M, N, K):
    """test_quantized.py:30-57, 275-330: |qmm(x, wq) - x @ dq(wq).T|_inf < 1e-3 in fp32."""
    rng = np.random.default_rng(1)
    x = rng.standard_normal((M, K)).astype(np.float32) / np.sqrt(K)
    w = rng.standard_normal((N, K)).astype(np.float32)
    wq, s, b = orc.quantize(w, gs, bits, dtype="f32")
    w_hat = orc.dequantize(wq, s, b, gs, bits, t_bf16=False)
    y = orc.quantized_matmul(x, wq, s, b, True, gs, bits, mode="A")
    assert y.shape == (M, N)
    assert np.abs(y - x @ w_hat.T).max() < 1e-3


def test_throw():
    """test_quantized.py:223-237."""
    x = np.ones((3, 512), np.float32)
    w = np.ones((512, 512), np.float32)
    wq, s, b = orc.quantize(w, 64, 2, dtype="f32")
    with pytest.raises(ValueError):
        orc.quantized_matmul(x, wq.T, s, b)
    with pytest.raises(ValueError):
        orc.quantized_matmul(x, wq, s.T, b)
    with pytest.raises(ValueError):
        orc.quantized_matmul(x, wq, s, b, True, 32, 2)
    with pytest.raises(ValueError):
        orc.quantize(np.ones((4, 100), np.float32), 64, 4)
    assert orc.quantized_matmul(x, wq, s, b, True, 64, 2).shape == (3, 512)


# ---------------------------------------------------------------- fast ops vs their pure-op statement (test_fast.py)
def _rope_orig(x, dims, traditional, base, scale, offset, freqs=None):
    """test_fast.py:10-47 in numpy/float64."""
    N = x.shape[-2] + offset
    half = dims // 2
    pos = np.arange(offset, N, dtype=np.float64) * scale
    inv = np.exp(-np.arange(half, dtype=np.float64) * (np.log(base) / half)) if freqs is None else 1.0 / freqs.astype(np.float64)
    th = pos[:, None] * inv[None]
    c, s = np.cos(th), np.sin(th)
    x = x.astype(np.float64)
    if traditional:
        x1, x2 = x[..., :dims:2], x[..., 1:dims:2]
        out = x.copy()
        out[..., :dims:2] = x1 * c - x2 * s
        out[..., 1:dims:2] = x1 * s + x2 * c
        return out
    x1, x2 = x[..., :half], x[..., half:dims]
    return np.concatenate([x1 * c - x2 * s, x1 * s + x2 * c, x[..., dims:]], axis=-1)


@pytest.mark.parametrize("traditional", [False, True])
@pytest.mark.parametrize("offset", [0, 3, 2047])
def test_rope_with_freqs(traditional, offset):
    """test_fast.py:169-251 (fp32 tolerance 1e-5 there; positions up to 2k here so 2e-4 on |x| <= ~4)."""
    rng = np.random.default_rng(2)
    x = rng.standard_normal((2, 4, 5, 64)).astype(np.float32)
    freqs = orc.llama3_freqs(64, 500000.0, 8.0, 1.0, 4.0, 131072)
    got = orc.rope(x, 64, traditional, None, 1.0, offset, freqs, mode="A")
    want = _rope_orig(x, 64, traditional, None, 1.0, offset, freqs)
    assert np.abs(got - want).max() < 2e-4
    got = orc.rope(x, 64, traditional, 10000.0, 1.0, offset, None, mode="A")
    want = _rope_orig(x, 64, traditional, 10000.0, 1.0, offset)
    assert np.abs(got - want).max() < 2e-4


def test_llama3_freqs_matches_python_formula():
    """PIE/models/llama/utils.py:22-39 evaluated in float64 with the SAME quirk (global == max == 131072)."""
    dims, base, factor, lo, hi, L = 128, 500000.0, 8.0, 1.0, 4.0, 131072
    f = base ** (np.arange(0, dims, 2, dtype=np.float64) / dims)
    wl = 2 * np.pi * f
    f2 = np.where(wl > L / lo, f * factor, f)
    med = (wl > L / hi) & (wl < L / lo)
    smooth = (L / wl - lo) / (hi - lo)
    want = np.where(med, f2 / ((1 - smooth) / factor + smooth), f2)
    got = orc.llama3_freqs(dims, base, factor, lo, hi, L)
    assert np.allclose(got, want, rtol=2e-6)


@pytest.mark.parametrize("dims", [31, 32, 33, 4099])
def test_rms_norm(dims):
    """test_fast.py:280-338."""
    rng = np.random.default_rng(3)
    x = rng.standard_normal((2, 3, dims)).astype(np.float32)
    w = rng.standard_normal(dims).astype(np.float32)
    want = x.astype(np.float64) / np.sqrt((x.astype(np.float64) ** 2).mean(-1, keepdims=True) + 1e-5) * w
    assert np.abs(orc.rms_norm(x, w, 1e-5, mode="A") - want).max() < 1e-5
    assert np.abs(orc.rms_norm(x, None, 1e-5, mode="A") * w - want).max() < 1e-5
    # mode B rounds to bf16 BEFORE the weight multiply (fast.cpp:104-108)
    xb, wb = orc.round_bf16(x), orc.round_bf16(w)
    n = orc.round_bf16(xb.astype(np.float64) / np.sqrt((xb.astype(np.float64) ** 2).mean(-1, keepdims=True) + 1e-5))
    gotb = orc.rms_norm(xb, wb, 1e-5, mode="B")
    assert np.mean(gotb == orc.round_bf16(n * wb)) > 0.995  # a few 1-ulp flips from fp32 vs fp64 reduction


@pytest.mark.parametrize("Hq,Hkv,L,T", [(8, 8, 1, 37), (8, 2, 1, 300), (4, 1, 5, 5), (8, 2, 7, 19)])
def test_sdpa_vs_naive(Hq, Hkv, L, T):
    """test_fast_sdpa.py:329-399: fast SDPA vs softmax(q k^T * scale + mask) v with GQA."""
    rng = np.random.default_rng(4)
    D = 64
    q = rng.standard_normal((2, Hq, L, D)).astype(np.float32)
    k = rng.standard_normal((2, Hkv, T, D)).astype(np.float32)
    v = rng.standard_normal((2, Hkv, T, D)).astype(np.float32)
    mask = orc.create_causal_mask(L, T - L) if L > 1 else None
    got = orc.scaled_dot_product_attention(q, k, v, D ** -0.5, mask, mode="A")
    rep = Hq // Hkv
    kk, vv = np.repeat(k, rep, 1).astype(np.float64), np.repeat(v, rep, 1).astype(np.float64)
    s = np.einsum("bhld,bhtd->bhlt", q.astype(np.float64), kk) * D ** -0.5
    if mask is not None:
        s = s + mask
    p = np.exp(s - s.max(-1, keepdims=True)); p /= p.sum(-1, keepdims=True)
    want = np.einsum("bhlt,bhtd->bhld", p, vv)
    assert np.abs(got - want).max() < 1e-5


def test_causal_mask_shape_and_values():
    m = orc.create_causal_mask(3, 2)  # PIE/models/base.py:18-34
    assert m.shape == (3, 5)
    assert (m[0] == [0, 0, 0, -1e9, -1e9]).all() and (m[2] == 0).all()


def test_kv_cache_semantics():
    """reusable.py:96-142 / cache.py:28-68: step-256 capacity, scalar offset, views of the first `offset` tokens."""
    c = orc.KVCache()
    k = np.ones((1, 2, 5, 8), np.float32)
    kk, vv = c.update_and_fetch(k, 2 * k)
    assert c.offset == 5 and kk.shape == (1, 2, 5, 8) and c.keys.shape[2] == 256
    kk, vv = c.update_and_fetch(3 * k[:, :, :1], k[:, :, :1])
    assert c.offset == 6 and (kk[:, :, 5] == 3).all() and (vv[:, :, :5] == 2).all()
    c.update_and_fetch(np.ones((1, 2, 300, 8), np.float32), np.ones((1, 2, 300, 8), np.float32))
    assert c.offset == 306 and c.keys.shape[2] == 512


def test_model_modes_agree_and_greedy_is_stable():
    """Tiny random-init model: mode B (bf16 op boundaries) stays within 2e-2 of mode A (fp32) on the logits, and the
    decode loop (prefill + L == 1 steps through the cache) equals a single full-sequence forward."""
    rng = np.random.default_rng(5)
    a = orc.ModelArgs(hidden_size=128, num_hidden_layers=2, intermediate_size=256, num_attention_heads=4,
                      num_key_value_heads=2, head_dim=32, vocab_size=320, rms_norm_eps=1e-5, rope_theta=500000.0,
                      max_position_embeddings=131072, tie_word_embeddings=True,
                      rope_scaling={"factor": 32.0, "low_freq_factor": 1.0, "high_freq_factor": 4.0})
    w = {}

    def put(name, n, k, scale):
        wq, s, b = orc.quantize(rng.standard_normal((n, k)).astype(np.float32) * scale, 64, 4)
        w[name + ".weight"], w[name + ".scales"], w[name + ".biases"] = wq, s, b

    put("model.embed_tokens", 320, 128, 0.1)
    for i in range(2):
        p = f"model.layers.{i}"
        for nm, n, k in (("self_attn.q_proj", 128, 128), ("self_attn.k_proj", 64, 128), ("self_attn.v_proj", 64, 128),
                         ("self_attn.o_proj", 128, 128), ("mlp.gate_proj", 256, 128), ("mlp.up_proj", 256, 128),
                         ("mlp.down_proj", 128, 256)):
            put(f"{p}.{nm}", n, k, k ** -0.5)
        w[p + ".input_layernorm.weight"] = orc.round_bf16(1 + 0.1 * rng.standard_normal(128).astype(np.float32))
        w[p + ".post_attention_layernorm.weight"] = orc.round_bf16(1 + 0.1 * rng.standard_normal(128).astype(np.float32))
    w["model.norm.weight"] = orc.round_bf16(1 + 0.1 * rng.standard_normal(128).astype(np.float32))
    ids = rng.integers(0, 320, (1, 9)).astype(np.int32)
    oa, ob = orc.LlamaOracle(a, w, mode="A"), orc.LlamaOracle(a, w, mode="B")
    la, lb = oa(ids, oa.make_cache()), ob(ids, ob.make_cache())
    assert np.linalg.norm(la - lb) / np.linalg.norm(la) < 2e-2
    cache = oa.make_cache()
    l1 = oa(ids[:, :6], cache)
    steps = [oa(ids[:, i:i + 1], cache)[:, 0] for i in range(6, 9)]
    inc = np.concatenate([l1, np.stack(steps, 1)], 1)
    assert np.abs(inc - la).max() < 1e-4 * np.abs(la).max() + 1e-5
    toks, _ = oa.generate_greedy(ids[0], 4)
    assert len(toks) == 4 and toks[0] == int(la[0, -1].argmax())
