"""TEST INFRASTRUCTURE ONLY -- CPU oracle for the quantized-Llama decode path.

Python face of ``oracle/pie_oracle.c`` (plain-C restatement of the reference algorithms) and of
``oracle/_ref/libpie_ref.so`` (the reference's own ``quantized.cpp`` compiled where it lies).
Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module.  The product package never does.

Citations: MLX = /root/reference/src/pie_core/external/mlx, PIE = /root/reference/src/proxy_inference_engine.

Tensors are numpy float32; a "bf16 tensor" is float32 with bf16-representable values.
``mode='A'``: every op in float32 (exactly what MLX's CPU backend computes for float32 inputs).
``mode='B'``: fp32 inside each op, bf16 rounding at every MLX op boundary (what the Metal
kernels compute for a bf16 checkpoint) -- the parity target
of the CUDA path.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass, field

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libpie_oracle.so")
_REF = os.path.join(_HERE, "_ref", "libpie_ref.so")

_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")
_u16p = np.ctypeslib.ndpointer(np.uint16, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")


def build(force: bool = False) -> None:
    """Compile the restatement (always possible: gcc) and, if /root/reference is mounted, _ref."""
    if force or not os.path.exists(_LIB) or (
        os.path.getmtime(_LIB) < os.path.getmtime(os.path.join(_HERE, "pie_oracle.c"))
    ):
        subprocess.check_call(["make", "-s", "-C", _HERE, os.path.join(_HERE, "libpie_oracle.so")])
    if os.path.isdir("/root/reference/src/pie_core/external/mlx") and (
        force or not os.path.exists(_REF)
        or os.path.getmtime(_REF) < os.path.getmtime(os.path.join(_HERE, "ref_harness.cpp"))
    ):
        subprocess.check_call(["make", "-s", "-C", _HERE, "ref"])


_lib = None
_ref = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB)
        L.orc_quantize.argtypes = [_f32p, _u32p, _f32p, _f32p, C.c_int, C.c_int, C.c_size_t, C.c_int]
        L.orc_unpack.argtypes = [_u32p, _u8p, C.c_size_t, C.c_int, C.c_int]
        L.orc_dequantize_rows.argtypes = [_u32p, _f32p, _f32p, C.c_void_p, C.c_size_t, C.c_int,
                                          C.c_int, C.c_int, C.c_int, _f32p]
        for name in ("orc_qmm_t_seq", "orc_qmm_t"):
            getattr(L, name).argtypes = [_f32p, _f32p, _u32p, _f32p, _f32p, C.c_int, C.c_int,
                                         C.c_int, C.c_int, C.c_int, C.c_int]
        L.orc_rms_norm.argtypes = [_f32p, C.c_void_p, _f32p, C.c_size_t, C.c_int, C.c_float, C.c_int]
        L.orc_rope.argtypes = [_f32p, _f32p, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int,
                               C.c_float, C.c_int, C.c_void_p, C.c_float, C.c_int]
        L.orc_llama3_freqs.argtypes = [_f32p, C.c_int] + [C.c_float] * 6
        L.orc_sdpa.argtypes = [_f32p, _f32p, _f32p, _f32p, C.c_int, C.c_int, C.c_int, C.c_int,
                               C.c_int, C.c_int, C.c_size_t, C.c_float, C.c_void_p, C.c_int, C.c_int]
        L.orc_silu_mul.argtypes = [_f32p, _f32p, _f32p, C.c_size_t, C.c_int]
        L.orc_add.argtypes = [_f32p, _f32p, _f32p, C.c_size_t, C.c_int]
        L.orc_round_bf16.argtypes = [_f32p, C.c_size_t]
        L.orc_f32_to_bf16.argtypes = [_f32p, _u16p, C.c_size_t]
        L.orc_bf16_to_f32.argtypes = [_u16p, _f32p, C.c_size_t]
        L.orc_max_threads.restype = C.c_int
        L.orc_set_threads.argtypes = [C.c_int]
        _lib = L
    return _lib


def ref():
    """The compiled reference (None when oracle/_ref was never built and cannot be built)."""
    global _ref
    if _ref is None:
        build()
        if not os.path.exists(_REF):
            return None
        R = C.CDLL(_REF)
        R.pie_ref_quantize_f32.argtypes = [_f32p, _u32p, _f32p, _f32p, C.c_int, C.c_int, C.c_size_t]
        R.pie_ref_quantize_bf16.argtypes = [_u16p, _u32p, _u16p, _u16p, C.c_int, C.c_int, C.c_size_t]
        R.pie_ref_qmm_t_f32.argtypes = [_f32p, _f32p, _u32p, _f32p, _f32p] + [C.c_int] * 5
        R.pie_ref_qmm_t_bf16.argtypes = [_u16p, _u16p, _u32p, _u16p, _u16p] + [C.c_int] * 5
        R.pie_ref_qmm_t_f32_mt.argtypes = [_f32p, _f32p, _u32p, _f32p, _f32p] + [C.c_int] * 6
        R.pie_ref_qmm_t_bf16_mt.argtypes = [_u16p, _u16p, _u32p, _u16p, _u16p] + [C.c_int] * 6
        R.pie_ref_qmm_n_f32.argtypes = [_f32p, _f32p, _u32p, _f32p, _f32p] + [C.c_int] * 5
        R.pie_ref_qmm_n_bf16.argtypes = [_u16p, _u16p, _u32p, _u16p, _u16p] + [C.c_int] * 5
        R.pie_ref_simd_max_size_f32.restype = C.c_int
        R.pie_ref_max_threads.restype = C.c_int
        _ref = R
    return _ref


# --------------------------------------------------------------------------------------------
# dtype helpers
# --------------------------------------------------------------------------------------------
def f32(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.float32)


def round_bf16(a) -> np.ndarray:
    out = f32(a).copy()
    lib().orc_round_bf16(out.reshape(-1), out.size)
    return out


def to_bf16_bits(a) -> np.ndarray:
    a = f32(a)
    out = np.empty(a.shape, np.uint16)
    lib().orc_f32_to_bf16(a.reshape(-1), out.reshape(-1), a.size)
    return out


def from_bf16_bits(a) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.uint16)
    out = np.empty(a.shape, np.float32)
    lib().orc_bf16_to_f32(a.reshape(-1), out.reshape(-1), a.size)
    return out


# --------------------------------------------------------------------------------------------
# operators (signatures follow mlx.core / mlx.core.fast)
# --------------------------------------------------------------------------------------------
def _check_quant_args(group_size: int, bits: int):
    # MLX/mlx/fast.cpp:839-866
    if group_size not in (32, 64, 128):
        raise ValueError(f"[quantize] The requested group size {group_size} is not supported.")
    if bits not in (2, 4, 8):
        raise ValueError(f"[quantize] oracle supports bits in (2,4,8), got {bits}.")


def quantize(w, group_size: int = 64, bits: int = 4, dtype: str = "bf16"):
    """mx.quantize (MLX/mlx/ops.cpp:4023-4029 -> cpu/quantized.cpp:603-661).

    dtype='bf16': w is rounded to bf16 first (a bf16 checkpoint) and scales/biases are stored as
    bf16 values; dtype='f32': pure float32.  Returns (wq uint32 [..., K*bits/32], scales, biases).
    """
    _check_quant_args(group_size, bits)
    w = f32(w)
    if w.ndim < 2:
        raise ValueError("[quantize] The matrix to be quantized must have at least 2 dimension")
    if w.shape[-1] % group_size != 0:
        raise ValueError(
            f"[quantize] The last dimension of the matrix needs to be divisible by the quantization "
            f"group size {group_size}. However the provided matrix has shape {w.shape}")
    if dtype == "bf16":
        w = round_bf16(w)
    K = w.shape[-1]
    wq = np.empty(w.shape[:-1] + (K * bits // 32,), np.uint32)
    s = np.empty(w.shape[:-1] + (K // group_size,), np.float32)
    b = np.empty_like(s)
    lib().orc_quantize(w.reshape(-1), wq.reshape(-1), s.reshape(-1), b.reshape(-1), bits,
                       group_size, w.size, 1 if dtype == "bf16" else 0)
    return wq, s, b


def unpack(wq, bits: int = 4) -> np.ndarray:
    wq = np.ascontiguousarray(wq, dtype=np.uint32)
    K = wq.shape[-1] * 32 // bits
    rows = wq.size // wq.shape[-1]
    q = np.empty((rows, K), np.uint8)
    lib().orc_unpack(wq.reshape(-1), q.reshape(-1), rows, K, bits)
    return q.reshape(wq.shape[:-1] + (K,))


def dequantize(wq, scales, biases, group_size: int = 64, bits: int = 4, ids=None,
               t_bf16: bool = True) -> np.ndarray:
    """mx.dequantize (MLX/mlx/fast.cpp:917-1029), optional row gather (quantized.py:100-107)."""
    wq = np.ascontiguousarray(wq, dtype=np.uint32)
    scales, biases = f32(scales), f32(biases)
    if wq.shape[-1] * 32 // bits != scales.shape[-1] * group_size:
        raise ValueError("[dequantize] Shape of scales and biases does not match the matrix")
    K = wq.shape[-1] * 32 // bits
    if ids is None:
        rows = wq.size // wq.shape[-1]
        out = np.empty((rows, K), np.float32)
        lib().orc_dequantize_rows(wq.reshape(-1), scales.reshape(-1), biases.reshape(-1), None,
                                  rows, K, bits, group_size, int(t_bf16), out.reshape(-1))
        return out.reshape(wq.shape[:-1] + (K,))
    ids = np.ascontiguousarray(ids, dtype=np.int32)
    out = np.empty((ids.size, K), np.float32)
    lib().orc_dequantize_rows(wq.reshape(-1), scales.reshape(-1), biases.reshape(-1),
                              ids.ctypes.data_as(C.c_void_p), ids.size, K, bits, group_size,
                              int(t_bf16), out.reshape(-1))
    return out.reshape(ids.shape + (K,))


def quantized_matmul(x, wq, scales, biases, transpose: bool = True, group_size: int = 64,
                     bits: int = 4, mode: str = "B", sequential: bool = False) -> np.ndarray:
    """mx.quantized_matmul (MLX/mlx/ops.cpp:3984-4021; checks :74-137); transpose=False -> _quantized_matmul_n.

    mode 'A': float32 result; 'B': result rounded to bf16; 'T': reference-exact bf16 scalar path
    (`_qmm_t<bfloat16_t>`, every op rounded).  sequential=True uses the strict k-order of `_qmm_t`.
    """
    if not transpose:
        return _quantized_matmul_n(x, wq, scales, biases, group_size, bits, mode)
    wq = np.ascontiguousarray(wq, dtype=np.uint32)
    x, scales, biases = f32(x), f32(scales), f32(biases)
    if scales.shape != biases.shape:
        raise ValueError("[quantized_matmul] Scales and biases should have the same shape.")
    if wq.shape[-1] * 32 // bits != scales.shape[-1] * group_size:
        raise ValueError("[quantized_matmul] The shapes of the weight and scales are incompatible "
                         "based on bits and group_size.")
    N, K = wq.shape[0], wq.shape[-1] * 32 // bits
    if x.shape[-1] != K:
        raise ValueError(f"[quantized_matmul] Last dimension of first input with shape {x.shape} "
                         f"does not match the expanded quantized matrix ({K}, {N})")
    M = x.size // K
    y = np.empty((M, N), np.float32)
    args = (y.reshape(-1), x.reshape(-1), wq.reshape(-1), scales.reshape(-1), biases.reshape(-1),
            M, N, K, bits, group_size)
    if mode == "T":
        lib().orc_qmm_t_seq(*args, 1)
    elif sequential:
        lib().orc_qmm_t_seq(*args, 0)
        if mode == "B":
            y = round_bf16(y)
    else:
        lib().orc_qmm_t(*args, 1 if mode == "B" else 0)
    return y.reshape(x.shape[:-1] + (N,))


def _quantized_matmul_n(x, wq, scales, biases, group_size, bits, mode) -> np.ndarray:
    """transpose=False: y[m, n] = sum_k x[m, k] * (scales[k, n/gs] * q[k, n] + biases[k, n/gs]), w quantised along N
    (`_qmm`, MLX/mlx/backend/cpu/quantized.cpp:38-94: k outer, `result[n] += xi * (scale * w + bias)`).  Leading batch
    dimensions of x and w broadcast as in MLX.  mode 'A': float32 in the reference's order; 'B': that, rounded to bf16;
    'T': `_qmm<bfloat16_t>` -- every operation rounded to bf16."""
    wq = np.ascontiguousarray(wq, dtype=np.uint32)
    x, scales, biases = f32(x), f32(scales), f32(biases)
    if scales.shape != biases.shape:
        raise ValueError("[quantized_matmul] Scales and biases should have the same shape.")
    N = wq.shape[-1] * 32 // bits
    if N != scales.shape[-1] * group_size:
        raise ValueError("[quantized_matmul] The shapes of the weight and scales are incompatible "
                         "based on bits and group_size.")
    K = wq.shape[-2]
    if x.shape[-1] != K:
        raise ValueError(f"[quantized_matmul] Last dimension of first input with shape {x.shape} "
                         f"does not match the expanded quantized matrix ({K}, {N})")
    q = unpack(wq, bits).astype(np.float32)                               # [..., K, N]
    s = np.repeat(scales, group_size, axis=-1)
    b = np.repeat(biases, group_size, axis=-1)
    r = round_bf16 if mode == "T" else (lambda a: a)
    wf = r(r(s * q) + b)                                                   # scale * w + bias
    y = np.zeros(np.broadcast_shapes(x.shape[:-2], wf.shape[:-2]) + (x.shape[-2], N), np.float32)
    for k in range(K):
        y = r(y + r(x[..., :, k, None] * wf[..., None, k, :]))
    return round_bf16(y) if mode == "B" else y


def softmax(x, mode: str = "B") -> np.ndarray:
    """mx.softmax(axis=-1) (MLX/mlx/backend/cpu/softmax.cpp): max, exp(x - max), sum, divide, in float32."""
    x = f32(x)
    e = np.exp(x - x.max(axis=-1, keepdims=True), dtype=np.float32)
    y = e / e.sum(axis=-1, keepdims=True, dtype=np.float32)
    return round_bf16(y) if mode != "A" else y


def quantized_scaled_dot_product_attention(queries, q_keys, q_values, scale: float, mask=None, group_size: int = 64,
                                           bits: int = 8, mode: str = "B") -> np.ndarray:
    """PIE/models/base.py:56-89.  queries [B, Hq, L, D]; q_keys / q_values = (quants, scales, biases) [B, Hkv, T, .].
    mode 'A': float32 with no intermediate rounding; 'B': bf16 rounding at every MLX op boundary (queries * scale,
    scores, + mask, softmax, out) with float32 inside the matmuls; 'T': additionally the matmuls are the x86 scalar
    bf16 kernels (`_qmm_t` / `_qmm`, every operation rounded) -- what the reference computes on this CPU."""
    queries = f32(queries)
    B, Hq, L, D = queries.shape
    Hkv = q_keys[0].shape[-3]
    n_rep = Hq // Hkv
    r = (lambda a: a) if mode == "A" else round_bf16
    mm = "A" if mode == "A" else mode
    queries = r(queries * np.float32(scale))
    if n_rep > 1:
        queries = queries.reshape(B, Hkv, n_rep, L, D)
    out = np.empty((B, Hkv, n_rep, L, q_values[0].shape[-1] * 32 // bits), np.float32)
    for b_ in range(B):
        for h in range(Hkv):
            qk = tuple(np.asarray(t)[b_, h] for t in q_keys)
            qv = tuple(np.asarray(t)[b_, h] for t in q_values)
            x = queries[b_, h].reshape(n_rep * L, D) if n_rep > 1 else queries[b_, h]
            sc = quantized_matmul(x, *qk, transpose=True, group_size=group_size, bits=bits, mode=mm,
                                  sequential=(mode == "T"))
            if mask is not None:
                sc = r(sc.reshape(n_rep, L, -1) + f32(mask)).reshape(n_rep * L, -1)
            sc = softmax(sc, mode)
            out[b_, h] = quantized_matmul(sc, *qv, transpose=False, group_size=group_size, bits=bits,
                                          mode=mm).reshape(n_rep, L, -1)
    return out.reshape(B, Hq, L, -1)


def rms_norm(x, weight, eps: float, mode: str = "B") -> np.ndarray:
    """mx.fast.rms_norm (MLX/mlx/fast.cpp:56-123)."""
    x = f32(x)
    if weight is not None:
        weight = f32(weight)
        if weight.ndim != 1 or weight.size != x.shape[-1]:
            raise ValueError("[rms_norm] weight must have the same size as the last dimension of x")
    out = np.empty_like(x)
    wp = None if weight is None else weight.ctypes.data_as(C.c_void_p)
    lib().orc_rms_norm(x.reshape(-1), wp, out.reshape(-1), x.size // x.shape[-1], x.shape[-1],
                       eps, int(mode == "B"))
    return out


def llama3_freqs(dims, base, factor, low_freq_factor, high_freq_factor, max_embedding_length,
                 global_embedding_length=None) -> np.ndarray:
    """Llama3RoPE.__init__ (PIE/models/llama/utils.py:22-39)."""
    if global_embedding_length is None:
        global_embedding_length = max_embedding_length
    out = np.empty(dims // 2, np.float32)
    lib().orc_llama3_freqs(out, dims, base, factor, low_freq_factor, high_freq_factor,
                           float(max_embedding_length), float(global_embedding_length))
    return out


def rope(x, dims: int, traditional: bool = False, base=None, scale: float = 1.0, offset: int = 0,
         freqs=None, mode: str = "B") -> np.ndarray:
    """mx.fast.rope (MLX/mlx/fast.cpp:353-511).  x: [..., L, D]."""
    x = f32(x)
    if x.ndim < 3:
        raise ValueError("[rope] Input must have at least 3 dimensions")
    if (base is None) == (freqs is None):
        raise ValueError("[rope] Exactly one of base or freqs must have a value.")
    fp = None
    if freqs is not None:
        freqs = f32(freqs)
        if freqs.ndim != 1 or freqs.shape[0] != dims // 2:
            raise ValueError(f"[rope] freqs must be one dimensional with size {dims // 2}")
        fp = freqs.ctypes.data_as(C.c_void_p)
    L, Dt = x.shape[-2], x.shape[-1]
    out = np.empty_like(x)
    lib().orc_rope(x.reshape(-1), out.reshape(-1), x.size // (L * Dt), L, Dt, dims,
                   int(traditional), scale, int(offset), fp, 0.0 if base is None else float(base),
                   int(mode == "B"))
    return out


def scaled_dot_product_attention(q, k, v, scale: float, mask=None, causal_offset: int = -1,
                                 mode: str = "B") -> np.ndarray:
    """mx.fast.scaled_dot_product_attention (MLX/mlx/fast.cpp:566-783), q [B,Hq,L,D], k/v [B,Hkv,T,D]."""
    q, k, v = f32(q), f32(k), f32(v)
    for t in (q, k, v):
        if t.ndim != 4:
            raise ValueError("[scaled_dot_product_attention] inputs expected to be rank 4")
    B, Hq, L, D = q.shape
    Hkv, T = k.shape[1], k.shape[2]
    if Hq % Hkv != 0:
        raise ValueError("[scaled_dot_product_attention] n_heads must be a multiple of n_kv_heads")
    mp = None
    if mask is not None:
        mask = f32(np.broadcast_to(f32(mask), (L, T)))
        mp = mask.ctypes.data_as(C.c_void_p)
    out = np.empty_like(q)
    lib().orc_sdpa(q.reshape(-1), k.reshape(-1), v.reshape(-1), out.reshape(-1), B, Hq, Hkv, L, T,
                   D, T, scale, mp, causal_offset, int(mode == "B"))
    return out


def create_causal_mask(N: int, offset: int = 0) -> np.ndarray:
    """PIE/models/base.py:18-34 (window_size / lengths unused on the Llama path)."""
    rinds = np.arange(offset + N)[None]
    linds = np.arange(offset, offset + N)[:, None]
    return ((linds < rinds) * -1e9).astype(np.float32)


def silu_mul(g, u, mode: str = "B") -> np.ndarray:
    g, u = f32(g), f32(u)
    out = np.empty_like(g)
    lib().orc_silu_mul(g.reshape(-1), u.reshape(-1), out.reshape(-1), g.size, int(mode == "B"))
    return out


def add(a, b, mode: str = "B") -> np.ndarray:
    a, b = f32(a), f32(b)
    out = np.empty_like(a)
    lib().orc_add(a.reshape(-1), b.reshape(-1), out.reshape(-1), a.size, int(mode == "B"))
    return out


# --------------------------------------------------------------------------------------------
# model: the wiring of PIE/models/llama/language.py
# --------------------------------------------------------------------------------------------
# --------------------------------------------------------------------------------------------
# sampler (PIE/engine/inference_engine.py:268-271; PIE/samplers/*.py) -- numpy restatement
# --------------------------------------------------------------------------------------------
def logprobs_of(logits) -> np.ndarray:
    """logits_f32 - logsumexp(logits_f32, axis=-1, keepdims=True)   (inference_engine.py:268-269)"""
    x = np.asarray(logits, dtype=np.float32)
    m = x.max(axis=-1, keepdims=True)
    return x - (m + np.log(np.exp(x - m).sum(axis=-1, keepdims=True, dtype=np.float32)))


def _categorical(x, uniforms) -> np.ndarray:
    """mx.random.categorical = arg-max of logits + Gumbel noise; the noise is given as uniforms in (0, 1)."""
    g = -np.log(-np.log(np.asarray(uniforms, dtype=np.float32)))
    return np.argmax(np.where(np.isneginf(x), -np.inf, x + g), axis=-1).astype(np.int32)


def sample_reference(logprobs, uniforms, temp, top_p=0.0, min_p=0.0, min_tokens_to_keep=1, top_k=-1) -> np.ndarray:
    """make_sampler (PIE/samplers/__init__.py:11-45) for temp > 0, statement for statement, with the random draw taken
    from `uniforms` [rows, vocab] (uniforms[r, i] belongs to TOKEN i, whatever order a sampler sorts the tokens in)."""
    lp = np.asarray(logprobs, dtype=np.float32)
    u = np.asarray(uniforms, dtype=np.float32)
    if 0 < top_p < 1.0:                                   # top_p.py:8-37
        x = lp * np.float32(1 / temp)
        p = np.exp(x - x.max(-1, keepdims=True)); p = (p / p.sum(-1, keepdims=True, dtype=np.float32)).astype(np.float32)
        idx = np.argsort(p, axis=-1, kind="stable")
        sp = np.take_along_axis(p, idx, -1)
        cum = np.cumsum(sp, axis=-1, dtype=np.float32)
        top = np.where(cum > 1 - top_p, sp, 0)
        with np.errstate(divide="ignore"):
            tok = _categorical(np.log(top), np.take_along_axis(u, idx, -1))
        return np.take_along_axis(idx, tok[:, None].astype(np.int64), -1)[:, 0].astype(np.int32)
    if min_p != 0.0:                                      # min_p.py:8-67
        x = lp * np.float32(1 / temp)
        idx = np.argsort(-x, axis=-1, kind="stable")
        sx = np.take_along_axis(x, idx, -1)
        remove = sx < sx[:, 0:1] + np.float32(np.log(min_p))
        remove[..., :min_tokens_to_keep] = False
        tok = _categorical(np.where(remove, -np.inf, sx), np.take_along_axis(u, idx, -1))
        return np.take_along_axis(idx, tok[:, None].astype(np.int64), -1)[:, 0].astype(np.int32)
    if top_k > 0:                                         # top_k.py:8-35
        x = lp * np.float32(1 / temp)
        mask_idx = np.argpartition(-x, kth=top_k - 1, axis=-1)[..., top_k:]
        masked = x.copy()
        np.put_along_axis(masked, mask_idx, -np.inf, axis=-1)
        return _categorical(masked, u)
    return _categorical(lp * np.float32(1 / temp), u)     # categorical.py:6-8


@dataclass
class ModelArgs:
    """PIE/models/llama/language.py:13-29."""
    hidden_size: int
    num_hidden_layers: int
    intermediate_size: int
    num_attention_heads: int
    rms_norm_eps: float
    vocab_size: int
    head_dim: int | None = None
    max_position_embeddings: int | None = None
    num_key_value_heads: int | None = None
    rope_theta: float = 10000.0
    rope_traditional: bool = False
    rope_scaling: dict | None = None
    tie_word_embeddings: bool = True
    model_type: str = "llama"
    quantization: dict = field(default_factory=lambda: {"group_size": 64, "bits": 4})


class KVCache:
    """ReusableKVCache / KVCache semantics (PIE/cache/kv_cache/reusable.py:96-142, cache.py:28-68):
    layout [B, n_kv, cap, D], cap grows in steps of 256, scalar offset."""

    def __init__(self, step: int = 256):
        self.keys = None
        self.values = None
        self.offset = 0
        self.step = step

    def update_and_fetch(self, k, v):
        needed = k.shape[2]
        prev = self.offset
        if self.keys is None or prev + needed > self.keys.shape[2]:
            B, H, _, D = k.shape
            cap = ((prev + needed + self.step - 1) // self.step) * self.step
            nk = np.zeros((B, H, cap, D), np.float32)
            nv = np.zeros((B, H, cap, v.shape[3]), np.float32)
            if self.keys is not None:
                nk[:, :, :prev] = self.keys[:, :, :prev]
                nv[:, :, :prev] = self.values[:, :, :prev]
            self.keys, self.values = nk, nv
        self.keys[:, :, prev:prev + needed] = k
        self.values[:, :, prev:prev + needed] = v
        self.offset += needed
        return self.keys[:, :, :self.offset], self.values[:, :, :self.offset]


class QuantizedKVCache:
    """QuantizedKVCache (PIE/cache/kv_cache/quantized.py:8-108): (quants, scales, biases) triples, capacity in steps
    of 256, mx.quantize of the new rows (bf16 tensors)."""

    def __init__(self, group_size: int = 64, bits: int = 8, step: int = 256):
        self.keys = None
        self.values = None
        self.offset = 0
        self.step = step
        self.group_size = group_size
        self.bits = bits

    def update_and_fetch(self, k, v):
        B, H, n, D = k.shape
        prev = self.offset
        if self.keys is None or prev + n > self.keys[0].shape[2]:
            new_steps = (self.step + n - 1) // self.step * self.step
            cap = (prev if self.keys is not None else 0) + new_steps

            def init(dim, old):
                t = (np.zeros((B, H, cap, dim * self.bits // 32), np.uint32),
                     np.zeros((B, H, cap, dim // self.group_size), np.float32),
                     np.zeros((B, H, cap, dim // self.group_size), np.float32))
                if old is not None:
                    for a, o in zip(t, old):
                        a[:, :, :prev] = o[:, :, :prev]
                return t
            self.keys, self.values = init(D, self.keys), init(v.shape[-1], self.values)
        self.offset += n
        for dst, src in zip(self.keys, quantize(k, self.group_size, self.bits)):
            dst[:, :, prev:self.offset] = src
        for dst, src in zip(self.values, quantize(v, self.group_size, self.bits)):
            dst[:, :, prev:self.offset] = src
        cut = lambda t: t[:, :, :self.offset]
        return tuple(map(cut, self.keys)), tuple(map(cut, self.values))


class LlamaOracle:
    """Model.__call__ (PIE/models/llama/language.py:199-210) over MLX-layout quantized weights.

    ``weights``: dict of numpy arrays keyed as the checkpoint schema:
    ``model.layers.{i}.self_attn.q_proj.{weight,scales,biases}`` ... (uint32 / float32 holding
    bf16 values), ``model.embed_tokens.*``, ``model.norm.weight``, ``lm_head.*``.
    """

    def __init__(self, args: ModelArgs, weights: dict, mode: str = "B"):
        self.args = args
        self.w = weights
        self.mode = mode
        self.gs = args.quantization["group_size"]
        self.bits = args.quantization["bits"]
        self.n_heads = args.num_attention_heads
        self.n_kv = args.num_key_value_heads or args.num_attention_heads
        self.head_dim = args.head_dim or args.hidden_size // self.n_heads
        self.scale = self.head_dim ** -0.5
        rs = args.rope_scaling or {}
        max_len = args.max_position_embeddings or 8192
        # language.py:55-73 -- max_embedding_length == global_embedding_length == max_position_embeddings
        self.freqs = llama3_freqs(self.head_dim, args.rope_theta, float(rs.get("factor", 1.0)),
                                  float(rs.get("low_freq_factor", 1.0)),
                                  float(rs.get("high_freq_factor", 1.0)), max_len, max_len)
        self.layers = list(range(args.num_hidden_layers))

    def make_cache(self, kv_bits=None, kv_group_size: int = 64):
        if kv_bits is not None:
            return [QuantizedKVCache(kv_group_size, kv_bits) for _ in self.layers]
        return [KVCache() for _ in self.layers]

    def _linear(self, x, prefix):
        return quantized_matmul(x, self.w[prefix + ".weight"], self.w[prefix + ".scales"],
                                self.w[prefix + ".biases"], True, self.gs, self.bits, self.mode)

    def __call__(self, ids, cache=None, trace: dict | None = None, last_only: bool = False):
        ids = np.asarray(ids, dtype=np.int32)
        B, L = ids.shape
        a, m = self.args, self.mode
        tb = m == "B"
        h = dequantize(self.w["model.embed_tokens.weight"], self.w["model.embed_tokens.scales"],
                       self.w["model.embed_tokens.biases"], self.gs, self.bits, ids=ids, t_bf16=tb)
        if cache is None:
            cache = [None] * len(self.layers)
        offset0 = cache[0].offset if cache[0] is not None else 0
        # create_attention_mask (base.py:37-53): additive mask only when T > 1
        mask = create_causal_mask(L, offset0) if L > 1 else None
        if mask is not None and tb:
            mask = round_bf16(mask)
        if trace is not None:
            trace["embed"] = h.copy()
        for i in self.layers:
            p = f"model.layers.{i}"
            c = cache[i]
            x1 = rms_norm(h, self.w[p + ".input_layernorm.weight"], a.rms_norm_eps, m)
            q = self._linear(x1, p + ".self_attn.q_proj")
            k = self._linear(x1, p + ".self_attn.k_proj")
            v = self._linear(x1, p + ".self_attn.v_proj")
            q = q.reshape(B, L, self.n_heads, -1).transpose(0, 2, 1, 3)
            k = k.reshape(B, L, self.n_kv, -1).transpose(0, 2, 1, 3)
            v = v.reshape(B, L, self.n_kv, -1).transpose(0, 2, 1, 3)
            off = c.offset if c is not None else 0
            q = rope(q, self.head_dim, a.rope_traditional, None, 1.0, off, self.freqs, m)
            k = rope(k, self.head_dim, a.rope_traditional, None, 1.0, off, self.freqs, m)
            if c is not None:
                k, v = c.update_and_fetch(k, v)
            if isinstance(c, QuantizedKVCache):   # base.py:101-110
                o = quantized_scaled_dot_product_attention(q, k, v, self.scale, mask, c.group_size, c.bits, m)
            else:
                o = scaled_dot_product_attention(q, np.ascontiguousarray(k), np.ascontiguousarray(v),
                                                 self.scale, mask, -1, m)
            o = o.transpose(0, 2, 1, 3).reshape(B, L, -1)
            r = self._linear(o, p + ".self_attn.o_proj")
            h = add(h, r, m)
            x2 = rms_norm(h, self.w[p + ".post_attention_layernorm.weight"], a.rms_norm_eps, m)
            g = self._linear(x2, p + ".mlp.gate_proj")
            u = self._linear(x2, p + ".mlp.up_proj")
            act = silu_mul(g, u, m)
            r = self._linear(act, p + ".mlp.down_proj")
            h = add(h, r, m)
            if trace is not None:
                trace[f"layer{i}.attn_out"] = o.copy()
                trace[f"layer{i}.h"] = h.copy()
        if last_only:
            h = h[:, -1:, :]
        hn = rms_norm(h, self.w["model.norm.weight"], a.rms_norm_eps, m)
        if a.tie_word_embeddings:
            logits = self._linear(hn, "model.embed_tokens")  # as_linear, quantized.py:109-124
        else:
            logits = self._linear(hn, "lm_head")
        if trace is not None:
            trace["final_norm"] = hn.copy()
            trace["logits"] = logits.copy()
        return logits

    def generate_greedy(self, prompt_ids, n_steps: int, trace_logits: bool = False):
        """generate_step with temp == 0 (PIE/engine/inference_engine.py:228-297): prefill the whole
        prompt in one forward, then one token per forward; argmax of the last position."""
        cache = self.make_cache()
        ids = np.asarray(prompt_ids, dtype=np.int32)[None]
        logits = self(ids, cache, last_only=True)
        toks, all_logits = [], []
        for step in range(n_steps):
            nxt = int(np.argmax(logits[0, -1]))
            toks.append(nxt)
            if trace_logits:
                all_logits.append(logits[0, -1].copy())
            if step + 1 < n_steps:
                logits = self(np.array([[nxt]], np.int32), cache)
        return toks, all_logits
