"""GPU parity tests: decode / prefill attention, paged KV, and the fused GEMV forms, against the CPU oracle."""
import ctypes as C

import numpy as np
import pytest
import torch

from parity_util import n2t_bf16, n2t_words, rel_l2, t2n

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    from oracle import oracle as orc
    from proxy_inference_engine_b200 import ops
    return orc, ops


@pytest.mark.parametrize("B,Hq,Hkv,D,T", [(1, 4, 4, 64, 1), (2, 8, 2, 128, 7), (1, 32, 8, 128, 300),
                                           (2, 8, 1, 64, 1031), (1, 6, 3, 128, 2100), (3, 16, 2, 128, 65)])
def test_sdpa_decode(env, B, Hq, Hkv, D, T):
    """test_fast_sdpa.py:329-399 (sdpa_vector incl. GQA) with the cache-view layout of update_and_fetch."""
    orc, ops = env
    rng = np.random.default_rng(T + D)
    cap = ((T + 255) // 256) * 256
    q = orc.round_bf16(rng.standard_normal((B, Hq, 1, D)).astype(np.float32))
    k = orc.round_bf16(rng.standard_normal((B, Hkv, cap, D)).astype(np.float32))
    v = orc.round_bf16(rng.standard_normal((B, Hkv, cap, D)).astype(np.float32))
    kt, vt = n2t_bf16(k), n2t_bf16(v)
    got = t2n(ops.scaled_dot_product_attention(n2t_bf16(q), kt[:, :, :T], vt[:, :, :T], scale=D ** -0.5, mask=None))
    ref = orc.scaled_dot_product_attention(q, k[:, :, :T].copy(), v[:, :, :T].copy(), D ** -0.5, None, -1, "B")
    assert got.shape == ref.shape
    assert rel_l2(got, ref) < 3e-3, rel_l2(got, ref)


@pytest.mark.parametrize("nsplit", [1, 2, 5, 16])
def test_decode_attn_splits_and_ragged_positions(env, nsplit):
    orc, ops = env
    rng = np.random.default_rng(nsplit)
    B, Hq, Hkv, D, cap = 3, 8, 2, 128, 512
    q = orc.round_bf16(rng.standard_normal((B, Hq, D)).astype(np.float32))
    k = orc.round_bf16(rng.standard_normal((B, Hkv, cap, D)).astype(np.float32))
    v = orc.round_bf16(rng.standard_normal((B, Hkv, cap, D)).astype(np.float32))
    pos = np.array([0, 130, 511], np.int32)
    kt, vt = n2t_bf16(k), n2t_bf16(v)
    got = t2n(ops.decode_attention(n2t_bf16(q), kt, vt, ops.kv_layout(kt), torch.from_numpy(pos).cuda(), D ** -0.5,
                                   nsplit=nsplit)).reshape(B, Hq, D)
    for b in range(B):
        T = pos[b] + 1
        ref = orc.scaled_dot_product_attention(q[b:b + 1, :, None], k[b:b + 1, :, :T].copy(), v[b:b + 1, :, :T].copy(),
                                               D ** -0.5, None, -1, "B")[0, :, 0]
        assert rel_l2(got[b], ref) < 3e-3
    # replaying with the same workspace must give the same answer (counters self-clean)
    again = t2n(ops.decode_attention(n2t_bf16(q), kt, vt, ops.kv_layout(kt), torch.from_numpy(pos).cuda(), D ** -0.5,
                                     nsplit=nsplit)).reshape(B, Hq, D)
    assert np.array_equal(got, again)


def test_paged_attention_matches_contiguous(env):
    """No reference implementation exists for paged KV (stub, PIE/cache/kv_cache/paged.py): oracle = contiguous."""
    orc, ops = env
    from proxy_inference_engine_b200.cache import PagedKVCache
    rng = np.random.default_rng(3)
    B, Hq, Hkv, D, T = 2, 8, 2, 128, 200
    n_pg = 4
    table = torch.from_numpy(rng.permutation(B * n_pg).astype(np.int32).reshape(B, n_pg)).cuda()
    kp = torch.zeros((B * n_pg, 64, Hkv, D), dtype=torch.bfloat16, device="cuda")
    cache = PagedKVCache(kp, torch.zeros_like(kp), table)
    k = orc.round_bf16(rng.standard_normal((B, Hkv, T, D)).astype(np.float32))
    v = orc.round_bf16(rng.standard_normal((B, Hkv, T, D)).astype(np.float32))
    cache.update_and_fetch(n2t_bf16(k[:, :, :150]), n2t_bf16(v[:, :, :150]))   # ragged chunks across page edges
    cache.update_and_fetch(n2t_bf16(k[:, :, 150:]), n2t_bf16(v[:, :, 150:]))
    gk, gv = cache.gather()
    assert np.array_equal(t2n(gk), k) and np.array_equal(t2n(gv), v)
    q = orc.round_bf16(rng.standard_normal((B, Hq, D)).astype(np.float32))
    pos = torch.full((B,), T - 1, dtype=torch.int32, device="cuda")
    got = t2n(ops.decode_attention(n2t_bf16(q), cache.k_pool, cache.v_pool, cache.layout(), pos, D ** -0.5, nsplit=3))
    ref = orc.scaled_dot_product_attention(q[:, :, None], k, v, D ** -0.5, None, -1, "B")[:, :, 0]
    assert rel_l2(got.reshape(B, Hq, D), ref) < 3e-3


@pytest.mark.parametrize("B,Hq,Hkv,D,L,off", [(1, 4, 2, 64, 16, 0), (2, 8, 2, 128, 33, 0), (1, 8, 8, 128, 5, 40),
                                              (1, 4, 2, 128, 300, 0), (2, 4, 1, 128, 150, 77), (1, 2, 2, 64, 257, 63),
                                              (1, 8, 2, 128, 64, 64)])
def test_sdpa_prefill_causal(env, B, Hq, Hkv, D, L, off):
    orc, ops = env
    rng = np.random.default_rng(L + off)
    T = off + L
    q = orc.round_bf16(rng.standard_normal((B, Hq, L, D)).astype(np.float32))
    k = orc.round_bf16(rng.standard_normal((B, Hkv, T, D)).astype(np.float32))
    v = orc.round_bf16(rng.standard_normal((B, Hkv, T, D)).astype(np.float32))
    mask = ops.CausalMask(L, off)
    got = t2n(ops.scaled_dot_product_attention(n2t_bf16(q), n2t_bf16(k), n2t_bf16(v), scale=D ** -0.5, mask=mask))
    ref = orc.scaled_dot_product_attention(q, k, v, D ** -0.5, orc.create_causal_mask(L, off), -1, "B")
    assert got.shape == ref.shape and rel_l2(got, ref) < 3e-3


def _fused(ops, out, x, segs, M, K, bits, norm_w, epi, residual, ksplit=0):
    from proxy_inference_engine_b200 import _lib
    from proxy_inference_engine_b200.ops import _p, _stream
    L = _lib.load()
    arr = (_lib.QmvSeg * len(segs))()
    for i, (wq, s, b) in enumerate(segs):
        arr[i].wq, arr[i].scales, arr[i].biases, arr[i].N = wq.data_ptr(), s.data_ptr(), b.data_ptr(), wq.shape[0]
    ntot = segs[0][0].shape[0] if epi == 2 else sum(s[0].shape[0] for s in segs)
    nb = L.pie_qmv_workspace_bytes(M, ntot, K, bits, int(epi == 2), ksplit)
    ws = torch.zeros(max(nb, 256), dtype=torch.uint8, device="cuda")
    for _ in range(2):  # twice on the same workspace: counters must self-clean
        st = L.pie_qmv_fused(_p(out), out.shape[1], _p(x), x.shape[1], M, K, arr, len(segs), 64, bits, _p(norm_w), 1e-5,
                             epi, _p(residual), ksplit, _p(ws), ws.numel(), 1, _stream())
        _lib.check(st, "qmv_fused")
        if epi == 1:
            break  # in-place residual: one pass only
    torch.cuda.synchronize()


@pytest.mark.parametrize("bits", [4, 8])
@pytest.mark.parametrize("M,ksplit", [(1, 0), (3, 2), (8, 0)])
def test_fused_norm_qkv_resadd_swiglu(env, bits, M, ksplit):
    """One attention/MLP sub-block through the fused kernels vs the op-by-op oracle (language.py:126-154)."""
    orc, ops = env
    rng = np.random.default_rng(bits + M)
    h, f, nq, nkv = 512, 1024, 384, 128
    mk = lambda n, k: orc.quantize((rng.uniform(-1, 1, (n, k)) / np.sqrt(k)).astype(np.float32), 64, bits)
    up = lambda t: (n2t_words(t[0]), n2t_bf16(t[1]), n2t_bf16(t[2]))
    Wq, Wk, Wv, Wg, Wu, Wd = mk(nq, h), mk(nkv, h), mk(nkv, h), mk(f, h), mk(f, h), mk(h, f)
    x = orc.round_bf16(rng.standard_normal((M, h)).astype(np.float32) * 2)
    nw = orc.round_bf16(1 + 0.1 * rng.standard_normal(h).astype(np.float32))
    xn = orc.rms_norm(x, nw, 1e-5, "B")
    qm = lambda a, W: orc.quantized_matmul(a, W[0], W[1], W[2], True, 64, bits, mode="B")
    # norm + q|k|v
    out = torch.empty((M, nq + 2 * nkv), dtype=torch.bfloat16, device="cuda")
    _fused(ops, out, n2t_bf16(x), [up(Wq), up(Wk), up(Wv)], M, h, bits, n2t_bf16(nw), 0, None, ksplit)
    ref = np.concatenate([qm(xn, Wq), qm(xn, Wk), qm(xn, Wv)], axis=1)
    assert rel_l2(t2n(out), ref) < 4e-3
    # norm + SwiGLU
    act = torch.empty((M, f), dtype=torch.bfloat16, device="cuda")
    _fused(ops, act, n2t_bf16(x), [up(Wg), up(Wu)], M, h, bits, n2t_bf16(nw), 2, None, ksplit)
    ref_act = orc.silu_mul(qm(xn, Wg), qm(xn, Wu), "B")
    assert rel_l2(t2n(act), ref_act) < 6e-3
    # down + residual, in place
    res = n2t_bf16(x).clone()
    _fused(ops, res, n2t_bf16(ref_act), [up(Wd)], M, f, bits, None, 1, res, ksplit)
    ref_h = orc.add(x, qm(ref_act, Wd), "B")
    assert rel_l2(t2n(res), ref_h) < 3e-3


@pytest.mark.parametrize("bits", [4, 8])
@pytest.mark.parametrize("M", [9, 16, 32, 40])
def test_batched_decode_gemm_tcgen05(env, bits, M, monkeypatch):
    """9..32 rows per pass can go to the tcgen05 batched-decode GEMM (pie_qmd.cu): weights rounded once to bf16 like Metal
    qmm_t (kernels/quantized.h:954-1078).  Same sub-block as above, ragged N (320 rows: a partial 128-row tile), split-K
    for the narrow matrices, and M = 40 = one tcgen05 pass + one GEMV pass on the SAME workspace."""
    orc, ops = env
    monkeypatch.setenv("PIE_QMD", "2")  # everywhere it is supported (default: only where it measured faster)
    monkeypatch.setenv("PIE_QMI", "0")  # (8-bit weights would otherwise go to the int8 kernel: tests/test_gpu_qmi.py)
    rng = np.random.default_rng(100 + bits + M)
    h, f, nq, nkv = 1024, 2048, 320, 128
    mk = lambda n, k: orc.quantize((rng.uniform(-1, 1, (n, k)) / np.sqrt(k)).astype(np.float32), 64, bits)
    up = lambda t: (n2t_words(t[0]), n2t_bf16(t[1]), n2t_bf16(t[2]))
    Wq, Wk, Wv, Wg, Wu, Wd = mk(nq, h), mk(nkv, h), mk(nkv, h), mk(f, h), mk(f, h), mk(h, f)
    x = orc.round_bf16(rng.standard_normal((M, h)).astype(np.float32) * 2)
    nw = orc.round_bf16(1 + 0.1 * rng.standard_normal(h).astype(np.float32))
    xn = orc.rms_norm(x, nw, 1e-5, "B")

    def qm(a, W):  # bf16(weights), fp64 accumulate, one rounding
        w = orc.round_bf16(orc.dequantize(W[0], W[1], W[2], 64, bits, t_bf16=False)).astype(np.float64)
        return orc.round_bf16((a.astype(np.float64) @ w.T).astype(np.float32))

    out = torch.empty((M, nq + 2 * nkv), dtype=torch.bfloat16, device="cuda")
    _fused(ops, out, n2t_bf16(x), [up(Wq), up(Wk), up(Wv)], M, h, bits, n2t_bf16(nw), 0, None)
    ref = np.concatenate([qm(xn, Wq), qm(xn, Wk), qm(xn, Wv)], axis=1)
    assert rel_l2(t2n(out), ref) < 3e-3, rel_l2(t2n(out), ref)
    act = torch.empty((M, f), dtype=torch.bfloat16, device="cuda")
    _fused(ops, act, n2t_bf16(x), [up(Wg), up(Wu)], M, h, bits, n2t_bf16(nw), 2, None)
    ref_act = orc.silu_mul(qm(xn, Wg), qm(xn, Wu), "B")
    assert rel_l2(t2n(act), ref_act) < 6e-3, rel_l2(t2n(act), ref_act)
    res = n2t_bf16(x).clone()
    _fused(ops, res, n2t_bf16(ref_act), [up(Wd)], M, f, bits, None, 1, res)
    ref_h = orc.add(x, qm(ref_act, Wd), "B")
    assert rel_l2(t2n(res), ref_h) < 3e-3, rel_l2(t2n(res), ref_h)
    # plain operator form, and agreement with the GEMV (fp32 weights) within the bf16 rounding of the weights
    y = t2n(ops.quantized_matmul(n2t_bf16(x), *up(Wg), True, 64, bits))
    assert rel_l2(y, qm(x, Wg)) < 3e-3
    y8 = t2n(ops.quantized_matmul(n2t_bf16(x[:8]), *up(Wg), True, 64, bits))
    assert rel_l2(y[:8], y8) < 6e-3


def test_argmax_first_index_and_advance(env):
    orc, ops = env
    lg = torch.zeros((3, 1000), dtype=torch.bfloat16, device="cuda")
    lg[0, 17] = 2; lg[0, 900] = 2        # tie -> first index
    lg[1, 999] = 5
    lg[2] = -3; lg[2, 0] = -1
    pos = torch.tensor([4, 5, 6], dtype=torch.int32, device="cuda")
    seq = torch.zeros((3, 8), dtype=torch.int32, device="cuda")
    tok = ops.argmax(lg, positions_to_advance=pos, seq_out=seq)
    assert tok.tolist() == [17, 999, 0] and pos.tolist() == [5, 6, 7]
    assert seq[0, 5].item() == 17 and seq[1, 6].item() == 999 and seq[2, 7].item() == 0
