"""Operator surface of the path, with the names, argument meaning and error behaviour of the MLX operators
PIE's Llama model reaches: ``quantize``, ``dequantize``, ``quantized_matmul``, ``rms_norm``,
``rope``, ``scaled_dot_product_attention``.  Tensors are torch CUDA tensors; activations, scales, biases and
norm weights are bf16; packed weights are MLX ``uint32`` words held as ``torch.int32`` (same bits).

Every function is one or two calls into libpie_b200.so on the current CUDA stream.  No CPU fallback.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import KvLayout, KvqCache, QmvSeg

_BF16 = torch.bfloat16


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _as_words(w: torch.Tensor) -> torch.Tensor:
    if w.dtype == torch.int32:
        return w
    if w.dtype == torch.uint32:
        return w.view(torch.int32)
    raise ValueError(f"[quantized_matmul] The weight matrix should be uint32 but received {w.dtype}")


def _check_gs_bits(what, group_size, bits, full=False):
    # MLX/mlx/fast.cpp:839-866
    if group_size not in (32, 64, 128):
        raise ValueError(f"[{what}] The requested group size {group_size} is not supported. "
                         "The supported group sizes are 32, 64 and 128.")
    if bits not in (2, 3, 4, 6, 8):
        raise ValueError(f"[{what}] The requested number of bits {bits} is not supported. "
                         "The supported bits are 2, 3, 4, 6 and 8.")
    if bits in (3, 6) or (full and (group_size != 64 or bits not in (4, 8))):
        raise ValueError(f"[{what}] group_size={group_size}, bits={bits} is valid MLX but not built for sm_100a here "
                         "(supported: group_size 64, bits 4 or 8)")


class _Workspaces:
    """Zero-initialised scratch owned by the host side (the library never allocates)."""

    def __init__(self):
        self.bufs = {}

    def get(self, key, nbytes, device):
        buf = self.bufs.get((key, device))
        if buf is None or buf.numel() < nbytes:
            buf = torch.zeros(max(int(nbytes), 256), dtype=torch.uint8, device=device)
            self.bufs[(key, device)] = buf
        return buf


workspaces = _Workspaces()


def quantize(w: torch.Tensor, group_size: int = 64, bits: int = 4):
    """mx.quantize (MLX/mlx/ops.cpp:4023-4029; cpu/quantized.cpp:603-661).  Bit-exact with the reference."""
    _check_gs_bits("quantize", group_size, bits)
    if w.dim() < 2:
        raise ValueError("[quantize] The matrix to be quantized must have at least 2 dimension but it has only "
                         f"{w.dim()}.")
    if w.shape[-1] % group_size != 0:
        raise ValueError("[quantize] The last dimension of the matrix needs to be divisible by the quantization "
                         f"group size {group_size}. However the provided matrix has shape {tuple(w.shape)}")
    if w.dtype not in (_BF16, torch.float32):
        raise ValueError(f"[quantize] dtype {w.dtype} not built here (bfloat16 / float32)")
    w = w.contiguous()
    K = w.shape[-1]
    rows = w.numel() // K
    wq = torch.empty(w.shape[:-1] + (K * bits // 32,), dtype=torch.int32, device=w.device)
    s = torch.empty(w.shape[:-1] + (K // group_size,), dtype=w.dtype, device=w.device)
    b = torch.empty_like(s)
    L = _lib.load()
    _lib.check(L.pie_quantize(_p(w), 0 if w.dtype == _BF16 else 1, _p(wq), _p(s), _p(b), rows, K, group_size, bits,
                              _stream()), "quantize")
    return wq, s, b


def dequantize(w, scales, biases, group_size: int = 64, bits: int = 4, ids: torch.Tensor | None = None):
    """mx.dequantize (MLX/mlx/fast.cpp:917-1029), optionally gathering rows ``ids`` first
    (QuantizedEmbedding.__call__, MLX/python/mlx/nn/layers/quantized.py:100-107)."""
    _check_gs_bits("dequantize", group_size, bits)
    w = _as_words(w)
    if w.dim() < 2 or scales.dim() < 2 or biases.dim() < 2:
        raise ValueError("[dequantize] The matrix to be quantized must have at least 2 dimension.")
    if scales.shape != biases.shape:
        raise ValueError("[dequantize] Shape of scales and biases does not match")
    if w.shape[-1] * 32 // bits != scales.shape[-1] * group_size:
        raise ValueError(f"[dequantize] Shape of scales and biases does not match the matrix given the quantization "
                         f"parameters. Provided matrix of shape {tuple(w.shape)} and scales/biases of shape "
                         f"{tuple(scales.shape)} with group_size={group_size} and bits={bits}.")
    if scales.dtype != _BF16:
        raise ValueError("[dequantize] only bfloat16 scales/biases are built here")
    w, scales, biases = w.contiguous(), scales.contiguous(), biases.contiguous()
    K = w.shape[-1] * 32 // bits
    table_rows = w.numel() // w.shape[-1]
    if ids is None:
        out_shape, n_rows, idp = w.shape[:-1] + (K,), table_rows, None
    else:
        ids = ids.to(torch.int32).contiguous()
        out_shape, n_rows, idp = tuple(ids.shape) + (K,), ids.numel(), _p(ids)
    out = torch.empty(out_shape, dtype=_BF16, device=w.device)
    L = _lib.load()
    _lib.check(L.pie_dequantize(_p(out), _p(w), _p(scales), _p(biases), idp, n_rows, table_rows, K, group_size, bits,
                                _stream()), "dequantize")
    return out


def quantized_matmul(x, w, scales, biases, transpose: bool = True, group_size: int = 64, bits: int = 4):
    """mx.quantized_matmul (MLX/mlx/ops.cpp:3984-4021; validation :74-137), transpose=True only."""
    w = _as_words(w)
    if scales.shape != biases.shape:
        raise ValueError("[quantized_matmul] Scales and biases should have the same shape. Received scales with "
                         f"shape {tuple(scales.shape)} and biases with {tuple(biases.shape)}")
    if w.dim() < 2 or w.shape[:-1] != scales.shape[:-1]:
        raise ValueError("[quantized_matmul] Weight, scales and biases should have the same batch shape.")
    if w.shape[-1] * 32 // bits != scales.shape[-1] * group_size:
        raise ValueError("[quantized_matmul] The shapes of the weight and scales are incompatible based on bits and "
                         f"group_size. w.shape() == {tuple(w.shape)} and scales.shape() == {tuple(scales.shape)} "
                         f"with group_size={group_size} and bits={bits}")
    if not transpose:
        raise ValueError("[quantized_matmul] transpose=False is not on the QuantizedLinear path and is not built here")
    N, K = w.shape[-2], w.shape[-1] * 32 // bits
    if x.shape[-1] != K:
        raise ValueError(f"[quantized_matmul] Last dimension of first input with shape {tuple(x.shape)} does not "
                         f"match the expanded quantized matrix ({K}, {N}) computed from shape {tuple(w.shape)} with "
                         f"group_size={group_size}, bits={bits} and transpose=true")
    if not x.dtype.is_floating_point:
        raise ValueError("[quantized_matmul] Only real floating types are supported but x.dtype() == " f"{x.dtype}")
    _check_gs_bits("quantized_matmul", group_size, bits)
    if w.dim() != 2:
        raise ValueError("[quantized_matmul] batched weights are not built here")
    if x.dtype != _BF16 or scales.dtype != _BF16:
        raise ValueError("[quantized_matmul] only bfloat16 activations/scales/biases are built here")
    x2 = x.reshape(-1, K).contiguous()
    w, scales, biases = w.contiguous(), scales.contiguous(), biases.contiguous()
    M = x2.shape[0]
    y = torch.empty((M, N), dtype=_BF16, device=x.device)
    L = _lib.load()
    nb = L.pie_qmv_workspace_bytes(min(M, 32) if M else 1, N, K, bits, 0, 0) if bits in (4, 8) and group_size == 64 else 0
    ws = workspaces.get("qmv", nb, x.device)
    fn = L.pie_qmv if M <= 32 else L.pie_qmm
    _lib.check(fn(_p(y), _p(x2), _p(w), _p(scales), _p(biases), M, N, K, group_size, bits, _p(ws), ws.numel(),
                  _stream()), "quantized_matmul")
    return y.reshape(x.shape[:-1] + (N,))


def rms_norm(x, weight, eps: float):
    """mx.fast.rms_norm (MLX/mlx/fast.cpp:56-123): bf16(bf16(x * rsqrt(mean(x^2) + eps)) * weight)."""
    if x.dim() < 1:
        raise ValueError("[rms_norm] Input must have at least 1 dimension but got input with 0 dimensions.")
    if weight is not None:
        if weight.dim() != 1:
            raise ValueError(f"[rms_norm] weight must have 1 dimension but has {weight.dim()} dimensions.")
        if weight.shape[0] != x.shape[-1]:
            raise ValueError("[rms_norm] weight must have the same size as the last dimension of x but has "
                             f"{weight.shape[0]} elements.")
        if weight.dtype != _BF16:
            raise ValueError("[rms_norm] only bfloat16 is built here")
        weight = weight.contiguous()
    if x.dtype != _BF16:
        raise ValueError("[rms_norm] only bfloat16 is built here")
    x = x.contiguous()
    out = torch.empty_like(x)
    dims = x.shape[-1]
    L = _lib.load()
    _lib.check(L.pie_rms_norm(_p(out), _p(x), _p(weight), x.numel() // dims if dims else 0, dims, float(eps),
                              _stream()), "rms_norm")
    return out


def rope(x, dims: int, *, traditional: bool = False, base: float | None = None, scale: float = 1.0,
         offset: int = 0, freqs: torch.Tensor | None = None):
    """mx.fast.rope (MLX/mlx/fast.cpp:353-511).  x [..., L, D]; all leading dims are flattened."""
    if x.dim() < 3:
        raise ValueError(f"[rope] Input must have at least 3 dimensions but got input with {x.dim()} dimensions.")
    if (base is None) == (freqs is None):
        raise ValueError("[rope] Only one of base or freqs can have a value.")
    if freqs is not None:
        if freqs.dim() != 1:
            raise ValueError(f"[rope] freqs must be one dimensional but has {freqs.dim()} dimensions.")
        if freqs.shape[0] != dims // 2:
            raise ValueError(f"[rope] freqs must have the same size as dims / 2 = {dims // 2} but has size "
                             f"{freqs.shape[0]}.")
        freqs = freqs.to(torch.float32).contiguous()
    if x.dtype != _BF16:
        raise ValueError("[rope] only bfloat16 is built here")
    x = x.contiguous()
    out = torch.empty_like(x)
    Lq, D = x.shape[-2], x.shape[-1]
    L = _lib.load()
    _lib.check(L.pie_rope(_p(out), _p(x), x.numel() // (Lq * D), Lq, D, dims, int(traditional),
                          0.0 if base is None else float(base), float(scale), int(offset), _p(freqs), _stream()),
               "rope")
    return out


class CausalMask:
    """What create_attention_mask returns for L > 1 (PIE/models/base.py:37-53): the additive -1e9 causal mask
    [L, offset + L], kept symbolic because the attention kernel builds it from (offset, L)."""

    def __init__(self, L: int, offset: int):
        self.L, self.offset = L, offset
        self.shape = (L, offset + L)


def kv_layout(k: torch.Tensor, page_table=None, page_size: int = 0) -> KvLayout:
    """Strides (elements) of a [B, Hkv, T, D] view of a cache buffer, or of a page pool [pages, page, Hkv, D]."""
    lay = KvLayout()
    if page_table is None:
        if k.stride(3) != 1:
            raise ValueError("[scaled_dot_product_attention] keys/values must be contiguous in the feature dimension")
        lay.batch_stride, lay.head_stride, lay.tok_stride = k.stride(0), k.stride(1), k.stride(2)
        lay.page_table, lay.pt_stride, lay.page_size = None, 0, 0
    else:
        lay.batch_stride, lay.head_stride, lay.tok_stride = 0, k.stride(2), k.stride(1)
        lay.page_table, lay.pt_stride, lay.page_size = page_table.data_ptr(), page_table.stride(0), page_size
    return lay


def decode_attention(q, kbuf, vbuf, lay: KvLayout, positions, scale: float, nsplit: int = 0, max_context: int = 0):
    """Decode-step attention on explicit cache geometry.  q [B, Hq, D]; positions int32 [B] (context = pos + 1)."""
    B, Hq, D = q.shape
    Hkv = kbuf.shape[1] if lay.page_table is None else kbuf.shape[2]
    L = _lib.load()
    if nsplit <= 0:
        nsplit = L.pie_decode_attn_pick_splits(B, Hq, Hkv, max_context if max_context > 0 else 1)
    nb = L.pie_decode_attn_workspace_bytes(B, Hq, D, nsplit)
    # keyed by the size of the counter region: the counters must be zero at a launch, and a buffer that served another
    # (rows, heads) geometry holds that call's partial sums where this call's counters lie
    ws = workspaces.get(("attn", (B * Hq * 4 + 255) // 256), nb, q.device)
    out = torch.empty((B, Hq * D), dtype=_BF16, device=q.device)
    _lib.check(L.pie_decode_attn(_p(out), _p(q), _p(kbuf), _p(vbuf), C.byref(lay), _p(positions), B, Hq, Hkv, D,
                                 float(scale), nsplit, _p(ws), ws.numel(), _stream()), "scaled_dot_product_attention")
    return out


def scaled_dot_product_attention(q, k, v, *, scale: float, mask=None):
    """mx.fast.scaled_dot_product_attention (MLX/mlx/fast.cpp:566-783) for the two shapes the Llama path uses:
    L == 1 with mask None (decode) and L > 1 with the causal mask of create_attention_mask (prefill)."""
    for t in (q, k, v):
        if t.dim() != 4:
            raise ValueError(f"[scaled_dot_product_attention] input with shape {tuple(t.shape)} expected to be rank 4")
    B, Hq, Lq, D = q.shape
    if k.shape[0] != B or v.shape[0] != B:
        raise ValueError("[scaled_dot_product_attention] mismatching batch dimension for input")
    if k.shape[-1] != D:
        raise ValueError("[scaled_dot_product_attention] query, keys expected to have matching last dimension")
    Hkv, T = k.shape[1], k.shape[2]
    if v.shape[1] != Hkv:
        raise ValueError("[scaled_dot_product_attention] keys, values expected to have matching n_kv_heads")
    if Hq % Hkv != 0:
        raise ValueError("[scaled_dot_product_attention] n_heads must be a multiple of n_kv_heads")
    if q.dtype != _BF16 or k.dtype != _BF16 or v.dtype != _BF16:
        raise ValueError("[scaled_dot_product_attention] only bfloat16 is built here")
    if k.stride() != v.stride():
        raise ValueError("[scaled_dot_product_attention] keys and values must share one layout")
    lay = kv_layout(k)
    L = _lib.load()
    if Lq == 1 and mask is None:
        pos = torch.full((B,), T - 1, dtype=torch.int32, device=q.device)
        out = decode_attention(q.reshape(B, Hq, D).contiguous(), k, v, lay, pos, scale, max_context=T)
        return out.view(B, Hq, 1, D)
    if isinstance(mask, str) and mask == "causal":
        mask = CausalMask(Lq, T - Lq)
    if not isinstance(mask, CausalMask):
        raise ValueError("[scaled_dot_product_attention] only mask=None (L == 1), 'causal' or the CausalMask of "
                         "create_attention_mask are built here")
    if mask.offset + mask.L != T or mask.L != Lq:
        raise ValueError("[scaled_dot_product_attention] mask shape does not match the keys")
    out = torch.empty((B, Lq, Hq * D), dtype=_BF16, device=q.device)
    _lib.check(L.pie_prefill_attn(_p(out), _p(q.contiguous()), _p(k), _p(v), C.byref(lay), B, Hq, Hkv, Lq, D,
                                  mask.offset, float(scale), _stream()), "scaled_dot_product_attention")
    return out.view(B, Lq, Hq, D).permute(0, 2, 1, 3)


def prefill_attention(q, kbuf, vbuf, lay: KvLayout, offset: int, scale: float):
    """Causal attention of L > 1 new tokens over a cache that already holds them (T = offset + L) on explicit cache
    geometry, contiguous or paged: chunked prefill / prefix reuse.  q [B, Hq, L, D] -> [B, Hq, L, D]."""
    B, Hq, Lq, D = q.shape
    Hkv = kbuf.shape[1] if lay.page_table is None else kbuf.shape[2]
    out = torch.empty((B, Lq, Hq * D), dtype=_BF16, device=q.device)
    _lib.check(_lib.load().pie_prefill_attn(_p(out), _p(q.contiguous()), _p(kbuf), _p(vbuf), C.byref(lay), B, Hq, Hkv, Lq, D,
                                            int(offset), float(scale), _stream()), "scaled_dot_product_attention")
    return out.view(B, Lq, Hq, D).permute(0, 2, 1, 3)


def kvq_desc(keys, values, page_table=None, page_size: int = 0, group_size: int = 64, bits: int = 8) -> KvqCache:
    """pie_kvq_cache over the (quants, scales, biases) triples of a quantised KV cache: contiguous
    [B, Hkv, cap, D/4 | D/64] (QuantizedKVCache's own arrays) or pools [pages, Hkv, page_size, .] + a page table."""
    d = KvqCache()
    for name, t in zip(("k_quant", "k_scales", "k_biases", "v_quant", "v_scales", "v_biases"), (*keys, *values)):
        if not t.is_contiguous():
            raise ValueError("quantised KV cache buffers must be contiguous")
        setattr(d, name, t.data_ptr())
    d.capacity = keys[0].shape[2]
    if page_table is None:
        d.page_table, d.pt_stride, d.page_size = None, 0, 0
    else:
        d.page_table, d.pt_stride, d.page_size = page_table.data_ptr(), page_table.stride(0), page_size
    d.group_size, d.bits = group_size, bits
    return d


def kvq_append(qkv, desc: KvqCache, positions, freqs, Hq: int, Hkv: int, D: int, q_out=None):
    """Decode step: rotate q and k of the fused projection output [B, (Hq + 2 Hkv) D] at positions[b], quantise k and v
    into the cache (QuantizedKVCache.update_and_fetch, PIE/cache/kv_cache/quantized.py:92-106).  Returns q [B, Hq, D]."""
    B = qkv.shape[0]
    if q_out is None:
        q_out = torch.empty((B, Hq, D), dtype=_BF16, device=qkv.device)
    _lib.check(_lib.load().pie_kvq_append(_p(q_out), _p(qkv), C.byref(desc), _p(positions), _p(freqs), B, Hq, Hkv, D,
                                          _stream()), "kvq_append")
    return q_out


def kvq_write(desc: KvqCache, keys, values, offset: int):
    """Quantise k, v [B, Hkv, L, D] into cache rows offset .. offset + L - 1."""
    B, Hkv, L, D = keys.shape
    kc, vc = keys.contiguous(), values.contiguous()
    _lib.check(_lib.load().pie_kvq_write(C.byref(desc), _p(kc), _p(vc), B, Hkv, L, D, offset, _stream()), "kvq_write")


def decode_attention_q8(q, desc: KvqCache, positions, scale: float, Hkv: int, nsplit: int = 0, max_context: int = 0,
                        out=None):
    """quantized_scaled_dot_product_attention for one query token (PIE/models/base.py:56-89) over a quantised cache.
    q [B, Hq, D]; positions int32 [B] (context = pos + 1) -> [B, Hq * D]."""
    B, Hq, D = q.shape
    L = _lib.load()
    if nsplit <= 0:
        nsplit = L.pie_decode_attn_q8_pick_splits(B, Hq, Hkv, max_context if max_context > 0 else 1)
    nb = L.pie_decode_attn_workspace_bytes(B, Hq, D, nsplit)
    # keyed by the size of the counter region: the counters must be zero at a launch, and a buffer that served another
    # (rows, heads) geometry holds that call's partial sums where this call's counters lie
    ws = workspaces.get(("attn", (B * Hq * 4 + 255) // 256), nb, q.device)
    if out is None:
        out = torch.empty((B, Hq * D), dtype=_BF16, device=q.device)
    qc = q.contiguous()
    _lib.check(L.pie_decode_attn_q8(_p(out), _p(qc), C.byref(desc), _p(positions), B, Hq, Hkv, D, float(scale), nsplit,
                                    _p(ws), ws.numel(), _stream()), "decode_attn_q8")
    return out


def silu_mul(gate, up):
    """nn.silu(gate) * up with a bf16 rounding at each MLX op boundary (language.py:126-127)."""
    gate, up = gate.contiguous(), up.contiguous()
    out = torch.empty_like(gate)
    _lib.check(_lib.load().pie_silu_mul(_p(out), _p(gate), _p(up), gate.numel(), _stream()), "silu_mul")
    return out


def add(a, b):
    a, b = a.contiguous(), b.contiguous()
    out = torch.empty_like(a)
    _lib.check(_lib.load().pie_add(_p(out), _p(a), _p(b), a.numel(), _stream()), "add")
    return out


def argmax(logits, positions_to_advance=None, seq_out=None, out=None):
    """Greedy sampler (temp == 0): first arg-max of each row of bf16 logits [rows, vocab] -> int32 [rows]."""
    logits = logits.contiguous()
    rows, vocab = logits.shape
    if out is None:
        out = torch.empty((rows,), dtype=torch.int32, device=logits.device)
    _lib.check(_lib.load().pie_argmax(_p(out), _p(logits), rows, vocab, _p(positions_to_advance), _p(seq_out),
                                      0 if seq_out is None else seq_out.shape[1], _stream()), "argmax")
    return out


def new_rng_state(seed: int, device="cuda") -> torch.Tensor:
    """Device-resident Philox state of pie_sample: [seed, draw counter, scratch, scratch] (int64)."""
    return torch.tensor([int(seed) & 0x7FFFFFFFFFFFFFFF, 0, 0, 0], dtype=torch.int64, device=device)


def sample(logits, temp: float, top_p: float = 0.0, min_p: float = 0.0, min_tokens_to_keep: int = 1, top_k: int = -1,
           rng_state=None, uniforms=None, return_logprobs: bool = False, positions_to_advance=None, seq_out=None,
           out=None, logprobs_out=None):
    """The temperature branches of make_sampler (PIE/samplers/__init__.py:11-45) on bf16 logits [rows, vocab], in one
    launch together with logprobs = f32(logits) - logsumexp (inference_engine.py:268-269).  Returns tokens (int32 [rows])
    or (tokens, logprobs f32 [rows, vocab])."""
    if temp <= 0:
        raise ValueError("[sample] temp must be > 0 (temp == 0 is argmax)")
    if not (0 <= min_p <= 1.0):
        raise ValueError(f"`min_p` has to be a float in the [0, 1] interval, but is {min_p}")          # min_p.py:33-36
    if not isinstance(min_tokens_to_keep, int) or min_tokens_to_keep < 1:
        raise ValueError(f"`min_tokens_to_keep` has to be a positive integer, but is {min_tokens_to_keep}")  # :37-40
    logits = logits.contiguous()
    rows, vocab = logits.shape
    use_top_k = not (0 < top_p < 1.0) and min_p == 0.0 and top_k > 0
    if use_top_k and not (isinstance(top_k, int) and 0 < top_k < vocab):
        raise ValueError(f"`top_k` has to be an integer in the (0, {vocab}] interval, but is {top_k}.")  # top_k.py:23-27
    if logits.dtype != _BF16:
        raise ValueError("[sample] only bfloat16 logits are built here")
    if rng_state is None and uniforms is None:
        raise ValueError("[sample] needs rng_state (ops.new_rng_state) or uniforms")
    if out is None:
        out = torch.empty((rows,), dtype=torch.int32, device=logits.device)
    if return_logprobs and logprobs_out is None:
        logprobs_out = torch.empty((rows, vocab), dtype=torch.float32, device=logits.device)
    if uniforms is not None:
        uniforms = uniforms.to(torch.float32).contiguous()
    _lib.check(_lib.load().pie_sample(_p(out), _p(logits), rows, vocab, vocab, float(temp), float(top_p), float(min_p),
                                      int(min_tokens_to_keep), int(top_k), _p(rng_state), _p(uniforms), _p(logprobs_out),
                                      _p(positions_to_advance), _p(seq_out), 0 if seq_out is None else seq_out.shape[1],
                                      _stream()), "sample")
    return (out, logprobs_out) if return_logprobs else out


def launch_count() -> int:
    return int(_lib.load().pie_launch_count())
