"""8-bit weights at 9..32 activation rows on the integer tensor cores (csrc/pie_qmi.cu) against the oracle
(mx.quantized_matmul, MLX/mlx/ops.cpp:3984-4021; fp32 inside the op, bf16 at its boundary): one attention / MLP sub-block
with fused RMS norm, three row segments, residual add and SwiGLU, split-K and un-split, twice on the same workspace."""
import numpy as np
import pytest
import torch

from parity_util import n2t_bf16, n2t_words, rel_l2, t2n
from test_gpu_attn_fused import _fused

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    from oracle import oracle as orc
    from proxy_inference_engine_b200 import ops
    return orc, ops


@pytest.mark.parametrize("M", [9, 17, 32])
@pytest.mark.parametrize("h,f,nq,nkv", [(1024, 2048, 384, 128), (4096, 1536, 512, 256)])
def test_int8_tensor_core_form_matches_the_oracle(env, M, h, f, nq, nkv, monkeypatch):
    orc, ops = env
    from proxy_inference_engine_b200 import _lib
    monkeypatch.setenv("PIE_QMI", "1")
    rng = np.random.default_rng(M + h)
    mk = lambda n, k: orc.quantize((rng.uniform(-1, 1, (n, k)) / np.sqrt(k)).astype(np.float32), 64, 8)
    up = lambda t: (n2t_words(t[0]), n2t_bf16(t[1]), n2t_bf16(t[2]))
    Wq, Wk, Wv, Wg, Wu, Wd = mk(nq, h), mk(nkv, h), mk(nkv, h), mk(f, h), mk(f, h), mk(h, f)
    x = orc.round_bf16(rng.standard_normal((M, h)).astype(np.float32) * 2)
    x[min(3, M - 1)] *= 1e-3                      # rows of very different magnitude: the planes are scaled per row
    x[0, : h // 2] *= 2.0 ** -9                    # and a row whose small elements sit far below its maximum
    x = orc.round_bf16(x)
    nw = orc.round_bf16(1 + 0.1 * rng.standard_normal(h).astype(np.float32))
    xn = orc.rms_norm(x, nw, 1e-5, "B")
    qm = lambda a, W: orc.quantized_matmul(a, W[0], W[1], W[2], True, 64, 8, mode="B")

    n0 = _lib.load().pie_launch_count()
    out = torch.empty((M, nq + 2 * nkv), dtype=torch.bfloat16, device="cuda")
    _fused(ops, out, n2t_bf16(x), [up(Wq), up(Wk), up(Wv)], M, h, 8, n2t_bf16(nw), 0, None)
    assert _lib.load().pie_launch_count() - n0 == 4          # (prep + matmul) x 2 calls: the int8 form ran
    ref = np.concatenate([qm(xn, Wq), qm(xn, Wk), qm(xn, Wv)], axis=1)
    assert rel_l2(t2n(out), ref) < 2e-3, rel_l2(t2n(out), ref)
    assert np.abs(t2n(out) - ref).max() <= 2e-2 * np.abs(ref).max()
    act = torch.empty((M, f), dtype=torch.bfloat16, device="cuda")
    _fused(ops, act, n2t_bf16(x), [up(Wg), up(Wu)], M, h, 8, n2t_bf16(nw), 2, None)
    ref_act = orc.silu_mul(qm(xn, Wg), qm(xn, Wu), "B")
    assert rel_l2(t2n(act), ref_act) < 4e-3, rel_l2(t2n(act), ref_act)
    res = n2t_bf16(x).clone()
    _fused(ops, res, n2t_bf16(ref_act), [up(Wd)], M, f, 8, None, 1, res)
    ref_h = orc.add(x, qm(ref_act, Wd), "B")
    assert rel_l2(t2n(res), ref_h) < 2e-3, rel_l2(t2n(res), ref_h)
    # the plain operator, and agreement with the bf16 GEMV form of the same product
    y = t2n(ops.quantized_matmul(n2t_bf16(x), *up(Wg), True, 64, 8))
    assert rel_l2(y, qm(x, Wg)) < 2e-3
    monkeypatch.setenv("PIE_QMI", "0")
    y0 = t2n(ops.quantized_matmul(n2t_bf16(x), *up(Wg), True, 64, 8))
    assert rel_l2(y, y0) < 4e-3


def test_int8_form_zero_and_tiny_rows(env, monkeypatch):
    """An all-zero activation row (scale 0) and a denormal-sized one must come out as the bias-free zero / finite."""
    orc, ops = env
    monkeypatch.setenv("PIE_QMI", "1")
    rng = np.random.default_rng(5)
    W = orc.quantize((rng.uniform(-1, 1, (256, 512)) / 16).astype(np.float32), 64, 8)
    x = orc.round_bf16(rng.standard_normal((12, 512)).astype(np.float32))
    x[2] = 0.0
    x[5] *= 1e-30
    x = orc.round_bf16(x)
    y = t2n(ops.quantized_matmul(n2t_bf16(x), n2t_words(W[0]), n2t_bf16(W[1]), n2t_bf16(W[2]), True, 64, 8))
    ref = orc.quantized_matmul(x, W[0], W[1], W[2], True, 64, 8, mode="B")
    assert np.isfinite(y).all() and (y[2] == 0).all()
    assert rel_l2(np.delete(y, 5, 0), np.delete(ref, 5, 0)) < 2e-3
    assert np.abs(y[5]).max() <= 1e-25
