"""GPU parity tests of the operators, through the C ABI, against the CPU oracle (oracle/).  The cases restate
MLX's own property tests for the path (MLX/python/tests/test_quantized.py, test_fast.py, test_fast_sdpa.py)."""
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


def _rand_w(rng, n, k, scale=None):
    return (rng.uniform(-1, 1, (n, k)) * (scale or 1.0 / np.sqrt(k))).astype(np.float32)


@pytest.mark.parametrize("bits", [2, 4, 8])
@pytest.mark.parametrize("gs", [32, 64, 128])
def test_quantize_bit_exact(env, bits, gs):
    orc, ops = env
    rng = np.random.default_rng(bits * 100 + gs)
    w = orc.round_bf16(_rand_w(rng, 96, 512, 1.0))
    w[3] = 0.0                      # zeros round-trip exactly (test_quantized.py:24-28)
    w[5, :gs] = 1e-9                # scale clamp at eps
    wq_o, s_o, b_o = orc.quantize(w, gs, bits, dtype="bf16")
    wq, s, b = ops.quantize(n2t_bf16(w), gs, bits)
    assert np.array_equal(t2n(wq), wq_o)
    assert np.array_equal(t2n(s), s_o) and np.array_equal(t2n(b), b_o)
    # float32 flavour
    wq_o, s_o, b_o = orc.quantize(w, gs, bits, dtype="f32")
    wq, s, b = ops.quantize(torch.from_numpy(w).cuda(), gs, bits)
    assert np.array_equal(t2n(wq), wq_o) and np.array_equal(t2n(s), s_o) and np.array_equal(t2n(b), b_o)


@pytest.mark.parametrize("bits", [2, 4, 8])
def test_dequantize_exact_and_gather(env, bits):
    orc, ops = env
    rng = np.random.default_rng(bits)
    w = _rand_w(rng, 200, 256, 1.0)
    wq, s, b = orc.quantize(w, 64, bits)
    ref = orc.dequantize(wq, s, b, 64, bits)
    got = ops.dequantize(n2t_words(wq), n2t_bf16(s), n2t_bf16(b), 64, bits)
    assert np.array_equal(t2n(got), ref)
    assert np.abs(ref - orc.round_bf16(w)).max() <= np.abs(s).max() + 1e-2  # round-trip bound (test_quantized.py:11-23)
    ids = rng.integers(0, 200, (3, 5))
    got = ops.dequantize(n2t_words(wq), n2t_bf16(s), n2t_bf16(b), 64, bits, ids=torch.from_numpy(ids).cuda())
    assert np.array_equal(t2n(got), orc.dequantize(wq, s, b, 64, bits, ids=ids))


QMM_CASES = [(1, 128, 256), (2, 48, 512), (5, 33, 256), (8, 256, 1024), (9, 64, 768), (33, 128, 256),
             (1, 3, 256), (1, 4096, 4096), (4, 1024, 14336), (16, 528, 2048), (24, 96, 4096), (32, 256, 14336),
             (1, 6144, 4096), (2, 4720, 1024)]


@pytest.mark.parametrize("bits", [4, 8])
@pytest.mark.parametrize("M,N,K", QMM_CASES)
def test_quantized_matmul(env, bits, M, N, K):
    orc, ops = env
    rng = np.random.default_rng(M * 7 + N + K + bits)
    wq, s, b = orc.quantize(_rand_w(rng, N, K), 64, bits)
    x = orc.round_bf16(rng.standard_normal((M, K)).astype(np.float32))
    ref = orc.quantized_matmul(x, wq, s, b, True, 64, bits, mode="B")
    got = t2n(ops.quantized_matmul(n2t_bf16(x), n2t_words(wq), n2t_bf16(s), n2t_bf16(b), True, 64, bits))
    assert got.shape == ref.shape
    assert rel_l2(got, ref) < 3e-3, rel_l2(got, ref)
    # one bf16 ulp of slack on each element (fp32 reassociation may flip a rounding)
    assert np.all(np.abs(got - ref) <= np.abs(ref) * 2.0 ** -7 + 1e-3 * np.abs(ref).max())
    # against the dense product on dequantized weights in fp32 (test_quantized.py:30-57)
    dense = x @ orc.dequantize(wq, s, b, 64, bits, t_bf16=False).T
    assert rel_l2(got, dense) < 6e-3


@pytest.mark.parametrize("bits", [4, 8])
def test_split_k_workspace_serves_any_sequence_of_shapes(env, bits):
    """Regression: the split-K arrival counters sit in a FIXED region at the head of the caller's workspace.  With a
    shape-dependent split, a larger matrix run after a smaller one on the same (zeroed-once) workspace read the smaller
    one's partial sums as counters: 9..63 activation rows at Llama-8B widths were wrong from the second call on."""
    orc, ops = env
    for (M, N, K) in ((16, 1024, 4096), (16, 2048, 4096), (9, 4096, 4096), (32, 4096, 4096), (16, 4096, 2048),
                      (40, 6144, 4096), (32, 2384, 4096), (16, 4096, 14336), (33, 1024, 4096)):
        rng = np.random.default_rng(M + N + K)
        wq, s, b = orc.quantize(_rand_w(rng, N, K), 64, bits)
        x = orc.round_bf16(rng.standard_normal((M, K)).astype(np.float32))
        ref = orc.quantized_matmul(x, wq, s, b, True, 64, bits, mode="B")
        got = t2n(ops.quantized_matmul(n2t_bf16(x), n2t_words(wq), n2t_bf16(s), n2t_bf16(b), True, 64, bits))
        assert rel_l2(got, ref) < 3e-3, (M, N, K, rel_l2(got, ref))


# prefill sizes: M >= 64 runs on the tcgen05 / TMEM tile kernel (pie_qmm.cu): ragged M and N tiles, one and many K steps
QMM_TC_CASES = [(64, 128, 128), (70, 64, 512), (128, 128, 256), (200, 384, 512), (256, 200, 1024), (300, 1024, 4096),
                (1000, 264, 1152), (513, 4096, 2048), (96, 16, 14336)]


@pytest.mark.parametrize("bits", [4, 8])
@pytest.mark.parametrize("M,N,K", QMM_TC_CASES)
def test_quantized_matmul_prefill_tensor_core(env, bits, M, N, K):
    """Same contraction with the rounding structure of Metal qmm_t (kernels/quantized.h:954-1078): weights rounded once
    to bf16, fp32 accumulate.  Checked against that exact structure in fp32 on the host and against the fp32-dequant
    dense product of MLX's own test (test_quantized.py:119-169)."""
    orc, ops = env
    rng = np.random.default_rng(M + 3 * N + K + bits)
    wq, s, b = orc.quantize(_rand_w(rng, N, K), 64, bits)
    x = orc.round_bf16(rng.standard_normal((M, K)).astype(np.float32))
    got = t2n(ops.quantized_matmul(n2t_bf16(x), n2t_words(wq), n2t_bf16(s), n2t_bf16(b), True, 64, bits))
    assert got.shape == (M, N)
    w32 = orc.dequantize(wq, s, b, 64, bits, t_bf16=False)
    ref = orc.round_bf16(x.astype(np.float64) @ orc.round_bf16(w32).astype(np.float64).T)
    assert rel_l2(got, ref) < 2e-3, rel_l2(got, ref)
    assert np.all(np.abs(got - ref) <= np.abs(ref) * 2.0 ** -7 + 1e-3 * np.abs(ref).max())
    assert rel_l2(got, x @ w32.T) < 6e-3
    # and the decode GEMV on the same rows agrees within bf16 rounding of the weights
    rows = n2t_bf16(x[:8])
    small = t2n(ops.quantized_matmul(rows, n2t_words(wq), n2t_bf16(s), n2t_bf16(b), True, 64, bits))
    assert rel_l2(got[:8], small) < 6e-3


@pytest.mark.parametrize("gs,bits,K", [(32, 4, 96), (128, 8, 384), (64, 2, 192), (64, 4, 64)])
def test_quantized_matmul_generic_path(env, gs, bits, K):
    orc, ops = env
    rng = np.random.default_rng(K)
    wq, s, b = orc.quantize(_rand_w(rng, 40, K), gs, bits)
    x = orc.round_bf16(rng.standard_normal((3, K)).astype(np.float32))
    ref = orc.quantized_matmul(x, wq, s, b, True, gs, bits, mode="B")
    got = t2n(ops.quantized_matmul(n2t_bf16(x), n2t_words(wq), n2t_bf16(s), n2t_bf16(b), True, gs, bits))
    assert rel_l2(got, ref) < 3e-3


def test_quantized_matmul_throws(env):
    """MLX/python/tests/test_quantized.py:223-237 + ops.cpp:82-134."""
    orc, ops = env
    x = torch.ones((2, 256), dtype=torch.bfloat16, device="cuda")
    w = torch.ones((64, 256), dtype=torch.bfloat16, device="cuda")
    wq, s, b = ops.quantize(w, 64, 4)
    with pytest.raises(ValueError):
        ops.quantized_matmul(x, wq.float(), s, b)                       # weight not uint32
    with pytest.raises(ValueError):
        ops.quantized_matmul(x, wq, s[:, :2], b)                        # scales/biases mismatch
    with pytest.raises(ValueError):
        ops.quantized_matmul(x, wq, s, b, group_size=32)                # incompatible shapes
    with pytest.raises(ValueError):
        ops.quantized_matmul(x[:, :128], wq, s, b)                      # inner dimension
    with pytest.raises(ValueError):
        ops.quantized_matmul(x.to(torch.int32), wq, s, b)               # non-floating
    with pytest.raises(ValueError):
        ops.quantize(w, 48, 4)
    with pytest.raises(ValueError):
        ops.quantize(w, 64, 5)
    with pytest.raises(ValueError):
        ops.quantize(w[:, :100], 64, 4)
    assert ops.quantized_matmul(x[:0], wq, s, b).shape == (0, 64)      # empty input


@pytest.mark.parametrize("dims", [32, 33, 256, 4099])
def test_rms_norm(env, dims):
    orc, ops = env
    rng = np.random.default_rng(dims)
    x = orc.round_bf16(rng.standard_normal((3, 5, dims)).astype(np.float32) * 3)
    w = orc.round_bf16(1 + 0.1 * rng.standard_normal(dims).astype(np.float32))
    got = t2n(ops.rms_norm(n2t_bf16(x), n2t_bf16(w), 1e-5))
    ref = orc.rms_norm(x, w, 1e-5, mode="B")
    assert rel_l2(got, ref) < 2e-3
    assert np.mean(got != ref) < 0.02                                   # bf16 values agree except rare rounding flips
    with pytest.raises(ValueError):
        ops.rms_norm(n2t_bf16(x), n2t_bf16(w[:-1]), 1e-5)


@pytest.mark.parametrize("D,dims,trad,offset", [(64, 64, False, 0), (128, 128, False, 300), (128, 64, False, 7),
                                                 (64, 64, True, 2047), (128, 128, False, 100000)])
def test_rope(env, D, dims, trad, offset):
    orc, ops = env
    rng = np.random.default_rng(D + offset)
    x = orc.round_bf16(rng.standard_normal((2, 4, 9, D)).astype(np.float32))
    fr = orc.llama3_freqs(dims, 500000.0, 8.0, 1.0, 4.0, 131072)
    got = t2n(ops.rope(n2t_bf16(x), dims, traditional=trad, base=None, scale=1.0, offset=offset,
                       freqs=torch.from_numpy(fr).cuda()))
    ref = orc.rope(x, dims, trad, None, 1.0, offset, fr, mode="B")
    assert rel_l2(got, ref) < 2e-3
    if offset < 5000:  # base form: inv_freq comes from expf/logf, whose ulp differences scale with the position
        got_b = t2n(ops.rope(n2t_bf16(x), dims, traditional=trad, base=10000.0, scale=0.5, offset=offset))
        ref_b = orc.rope(x, dims, trad, 10000.0, 0.5, offset, None, mode="B")
        assert rel_l2(got_b, ref_b) < 5e-3
    with pytest.raises(ValueError):
        ops.rope(n2t_bf16(x), dims, base=None, freqs=None)
