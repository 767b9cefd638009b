"""Quantised KV cache on the GPU: pie_kvq_write / pie_kvq_append / pie_decode_attn_q8 through the C ABI
against tests/golden/ref_kvq.npz (outputs of the reference's own compiled quantize / _qmm_t / _qmm kernels,
tools/make_golden_kvq.py) and against the oracle's restatement of PIE/cache/kv_cache/quantized.py +
PIE/models/base.py:56-89.  Cache contents: bit-exact.  Attention output: 2e-2 of the float kernels (one bf16 rounding
at the store; fp32 inside, as the reference's Metal / Accelerate paths accumulate)."""
import os

import numpy as np
import pytest
import torch

from parity_util import make_oracle, n2t_bf16, rel_l2, t2n, tiny_args

pytestmark = pytest.mark.gpu

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_kvq.npz"))
CASES = [tuple(int(v) for v in c) for c in GOLD["cases"]]
TOL = 2e-2


def _filled_cache(ci, paged=False, seed=0):
    from proxy_inference_engine_b200.cache import PagedQuantizedKVCache, QuantizedKVCache
    g = lambda n: GOLD[f"c{ci}.{n}"]
    B, Hq, Hkv, T, D = CASES[ci]
    k, v = n2t_bf16(g("k")), n2t_bf16(g("v"))
    if paged:
        n_pg = (T + 63) // 64 + 1
        perm = torch.randperm(B * n_pg + 3, generator=torch.Generator().manual_seed(seed))[: B * n_pg]
        table = perm.reshape(B, n_pg).to(torch.int32).cuda()
        c = PagedQuantizedKVCache(B * n_pg + 3, Hkv, D, table)
    else:
        c = QuantizedKVCache(64, 8)
    cut = max(1, T // 2)
    c.update_and_fetch(k[:, :, :cut], v[:, :, :cut])
    if cut < T:
        c.update_and_fetch(k[:, :, cut:], v[:, :, cut:])
    return g, c


@pytest.mark.parametrize("paged", [False, True])
@pytest.mark.parametrize("ci", range(len(CASES)))
def test_cache_rows_bit_exact_with_reference_quantize(ci, paged):
    g, c = _filled_cache(ci, paged)
    B, Hq, Hkv, T, D = CASES[ci]
    assert c.offset == T
    qk, qv = c.gather() if paged else c.state
    for got, name in zip(qk + qv, ("kq", "ks", "kb", "vq", "vs", "vb")):
        want = g(name)
        got = got.contiguous()
        if want.dtype == np.uint16:
            got = got.view(torch.int16).cpu().numpy().view(np.uint16)
        else:
            got = t2n(got)
        assert np.array_equal(got, want), name


@pytest.mark.parametrize("nsplit", [1, 3, 0])
@pytest.mark.parametrize("paged", [False, True])
@pytest.mark.parametrize("ci", range(len(CASES)))
def test_decode_attention_matches_reference_kernels(ci, paged, nsplit):
    from proxy_inference_engine_b200 import ops
    g, c = _filled_cache(ci, paged)
    B, Hq, Hkv, T, D = CASES[ci]
    q = n2t_bf16(g("q")).reshape(B, Hq, D)
    pos = torch.full((B,), T - 1, dtype=torch.int32, device="cuda")
    out = ops.decode_attention_q8(q, c.desc(), pos, D ** -0.5, Hkv, nsplit=nsplit, max_context=T)
    got = out.float().cpu().numpy().reshape(B, Hq, 1, D)
    want = g("out_f32")
    assert np.abs(got - want).max() <= TOL * max(1.0, np.abs(want).max())
    assert rel_l2(got, want) < 5e-3


def test_ragged_contexts_and_split_counts_agree():
    """Per-row context lengths (positions[b]) and every split count give the same answer as the oracle."""
    from oracle import oracle as O
    from proxy_inference_engine_b200 import ops
    from proxy_inference_engine_b200.cache import QuantizedKVCache
    rng = np.random.default_rng(5)
    B, Hq, Hkv, T, D = 3, 8, 2, 700, 128
    k = O.round_bf16(rng.standard_normal((B, Hkv, T, D)).astype(np.float32))
    v = O.round_bf16(rng.standard_normal((B, Hkv, T, D)).astype(np.float32))
    q = O.round_bf16(rng.standard_normal((B, Hq, 1, D)).astype(np.float32))
    c = QuantizedKVCache(64, 8)
    c.update_and_fetch(n2t_bf16(k), n2t_bf16(v))
    ctx = [700, 1, 333]
    pos = torch.tensor([n - 1 for n in ctx], dtype=torch.int32, device="cuda")
    keys, vals = O.quantize(k, 64, 8), O.quantize(v, 64, 8)
    want = np.concatenate([
        O.quantized_scaled_dot_product_attention(q[b:b + 1], tuple(a[b:b + 1, :, :n] for a in keys),
                                                 tuple(a[b:b + 1, :, :n] for a in vals), D ** -0.5, None, 64, 8, "A")
        for b, n in enumerate(ctx)])
    for nsplit in (1, 2, 7, 16):
        out = ops.decode_attention_q8(n2t_bf16(q).reshape(B, Hq, D), c.desc(), pos, D ** -0.5, Hkv, nsplit=nsplit)
        got = out.float().cpu().numpy().reshape(B, Hq, 1, D)
        assert np.abs(got - want).max() <= TOL, nsplit


@pytest.mark.parametrize("D", [64, 128])
def test_append_rotates_and_quantises_like_rope_then_mx_quantize(D):
    """pie_kvq_append = rope(q), rope(k) -> bf16 -> mx.quantize into the row at positions[b]; v -> mx.quantize."""
    from oracle import oracle as O
    from proxy_inference_engine_b200 import ops
    from proxy_inference_engine_b200.cache import QuantizedKVCache
    from proxy_inference_engine_b200.llama import llama3_freqs
    B, Hq, Hkv = 3, 4, 2
    freqs = llama3_freqs(D, 500000.0, 8.0, 1.0, 4.0, 8192, 8192).cuda()
    g = torch.Generator().manual_seed(2)
    qkv = (torch.randn((B, (Hq + 2 * Hkv) * D), generator=g) * 3).to(torch.bfloat16).cuda()
    positions = torch.tensor([5, 0, 255], dtype=torch.int32, device="cuda")
    c = QuantizedKVCache(64, 8)
    c.reserve(B, Hkv, D, 256)
    q_out = ops.kvq_append(qkv, c.desc(), positions, freqs, Hq, Hkv, D)
    heads = qkv.view(B, Hq + 2 * Hkv, 1, D)
    for b in range(B):
        p = int(positions[b])
        q_ref = ops.rope(heads[b:b + 1, :Hq], D, base=None, offset=p, freqs=freqs)
        k_ref = ops.rope(heads[b:b + 1, Hq:Hq + Hkv], D, base=None, offset=p, freqs=freqs)
        assert torch.equal(q_out[b], q_ref[0, :, 0])
        for arrs, rows in ((c.keys, k_ref[0, :, 0]), (c.values, heads[b, Hq + Hkv:, 0])):
            wq, s, bb = O.quantize(t2n(rows), 64, 8)
            assert np.array_equal(t2n(arrs[0][b, :, p]), wq)
            assert np.array_equal(t2n(arrs[1][b, :, p]), s)
            assert np.array_equal(t2n(arrs[2][b, :, p]), bb)
    # untouched rows stay zero
    assert int(c.keys[0][1, :, 1:].abs().sum()) == 0


def test_model_with_quantized_cache_matches_oracle():
    """Model.__call__ with QuantizedKVCache: prefill + teacher-forced decode steps against the oracle running the same
    cache class (quantized.py) and base.py:56-89."""
    from proxy_inference_engine_b200.checkpoint import build_model, random_weights
    args = tiny_args(bits=4)
    w = random_weights(args, seed=1, device="cuda", gen_device="cpu")
    model, o = build_model(args, w), make_oracle(args, w, mode="B")
    ids = torch.randint(0, args.vocab_size, (2, 12), generator=torch.Generator().manual_seed(4))
    gc, oc = model.make_cache(kv_bits=8), o.make_cache(kv_bits=8)
    lg = model(ids.cuda().to(torch.int32), cache=gc)[:, -1].float().cpu().numpy()
    lo = o(ids.numpy(), oc, last_only=True)[:, -1]
    assert rel_l2(lg, lo) < TOL
    for step in range(6):
        tok = lo.argmax(-1).astype(np.int32)[:, None]
        lg = model(torch.from_numpy(tok).cuda(), cache=gc)[:, -1].float().cpu().numpy()
        lo = o(tok, oc, last_only=True)[:, -1]
        assert rel_l2(lg, lo) < TOL, step
    assert gc[0].offset == 18 and gc[0].keys[0].dtype == torch.int32


@pytest.mark.parametrize("head_dim,batch,paged,graph", [(64, 1, False, True), (128, 2, True, True), (64, 3, True, False)])
def test_decode_engine_with_quantized_cache(head_dim, batch, paged, graph):
    """DecodeEngine(kv_bits=8): the captured step (pie_kvq_append + pie_decode_attn_q8 per layer) against the oracle
    running QuantizedKVCache, teacher-forced over a context that crosses a 64-token page."""
    from parity_util import run_decode_parity
    r = run_decode_parity(tiny_args(bits=4, head_dim=head_dim), prompt_len=60, steps=8, batch=batch, paged=paged,
                          use_graph=graph, kv_bits=8)
    assert r["max_rel_err"] < TOL, r
    assert r["strong_mismatch"] == 0, r


def test_rejects_unsupported_configurations():
    from proxy_inference_engine_b200 import _lib, ops
    from proxy_inference_engine_b200.cache import QuantizedKVCache
    with pytest.raises(ValueError):
        QuantizedKVCache(32, 8)
    with pytest.raises(ValueError):
        QuantizedKVCache(64, 4)
    c = QuantizedKVCache(64, 8)
    c.reserve(1, 1, 128, 256)
    d = c.desc()
    d.bits = 4
    q = torch.zeros((1, 1, 128), dtype=torch.bfloat16, device="cuda")
    pos = torch.zeros(1, dtype=torch.int32, device="cuda")
    with pytest.raises(ValueError, match="UNSUPPORTED"):
        ops.decode_attention_q8(q, d, pos, 1.0, 1)
    d = c.desc()
    d.capacity = 255   # odd capacities are refused (scales are fetched in pairs of slots)
    with pytest.raises(ValueError, match="SHAPE"):
        ops.decode_attention_q8(q, d, pos, 1.0, 1)
