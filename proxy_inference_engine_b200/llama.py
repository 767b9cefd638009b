"""The Llama model surface of PIE (PIE/models/llama/language.py:13-223, utils.py:5-50, PIE/models/base.py:18-113),
class for class, over the B200 operators of ``ops`` / ``nn``.  ``Model.__call__(ids, cache)`` is the reference's
op-by-op forward (one library call per MLX op); ``decode.DecodeEngine`` runs the same step through the fused
kernels under a CUDA graph.
"""
from __future__ import annotations

import inspect
import math
from dataclasses import dataclass, field

import torch

from . import nn, ops
from .cache import KVCache, PagedKVCache, PagedQuantizedKVCache, QuantizedKVCache
from .tp import TPInfo, all_gather_last, all_reduce_sum


@dataclass
class ModelArgs:
    """language.py:13-29 (+ the ``quantization`` block of config.json, PIE/models/utils.py:95-111)."""
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
    lm_head_rows: int | None = None  # tensor parallel: vocabulary rows of the head this rank owns (tp.shard_args)

    @classmethod
    def from_dict(cls, params: dict):
        return cls(**{k: v for k, v in params.items() if k in inspect.signature(cls).parameters})


def llama3_freqs(dims, base, factor, low_freq_factor, high_freq_factor, max_embedding_length,
                 global_embedding_length) -> torch.Tensor:
    """Llama3RoPE.__init__ (utils.py:22-39) in float32, as mx evaluates it.  Wavelength-like: theta = pos / f."""
    f32 = torch.float32
    low_freq_wavelen = global_embedding_length / low_freq_factor
    high_freq_wavelen = global_embedding_length / high_freq_factor
    freqs = torch.tensor(base, dtype=f32) ** (torch.arange(0, dims, 2, dtype=f32) / dims)
    wavelens = 2 * math.pi * freqs
    freqs = torch.where(wavelens > low_freq_wavelen, freqs * factor, freqs)
    is_medium = (wavelens > high_freq_wavelen) & (wavelens < low_freq_wavelen)
    smooth = (max_embedding_length / wavelens - low_freq_factor) / (high_freq_factor - low_freq_factor)
    smooth_freqs = freqs / ((1 - smooth) / factor + smooth)
    return torch.where(is_medium, smooth_freqs, freqs).to(f32)


class Llama3RoPE(nn.Module):
    def __init__(self, max_embedding_length, global_embedding_length, dimensions, base, factor, low_freq_factor,
                 high_freq_factor, traditional=False, device="cuda"):
        self.dimensions = dimensions
        self.max_position_embeddings = max_embedding_length
        self.traditional = traditional
        self._freqs = llama3_freqs(dimensions, base, factor, low_freq_factor, high_freq_factor, max_embedding_length,
                                   global_embedding_length).to(device)

    def __call__(self, x, offset: int = 0):
        return ops.rope(x, self.dimensions, traditional=self.traditional, base=None, scale=1.0, offset=offset,
                        freqs=self._freqs)


def create_attention_mask(h, cache=None):
    """base.py:37-53: no mask for a single query token, else the causal mask offset by cache[0].offset."""
    T = h.shape[1]
    if T > 1:
        offset = cache[0].offset if cache is not None and cache[0] is not None else 0
        return ops.CausalMask(T, offset)
    return None


def quantized_scaled_dot_product_attention(queries, q_keys, q_values, scale, mask, group_size=64, bits=8, cache=None):
    """base.py:56-89.  One query token: pie_decode_attn_q8 over the quantised cache in place (K, V and the scores are
    never materialised).  L > 1 (prefill): the cached rows are expanded with mx.dequantize's arithmetic and go through
    the causal prefill kernel -- the same contraction as the reference's two quantized_matmul calls."""
    B, Hq, L, D = queries.shape
    if L == 1 and mask is None and cache is not None:
        Hkv = cache.keys[0].shape[1]
        pos = torch.full((B,), cache.offset - 1, dtype=torch.int32, device=queries.device)
        out = ops.decode_attention_q8(queries.reshape(B, Hq, D), cache.desc(), pos, scale, Hkv, max_context=cache.offset)
        return out.view(B, Hq, 1, D)
    if isinstance(q_keys, PagedQuantizedKVCache):
        q_keys, q_values = q_keys.gather()
    deq = lambda t: ops.dequantize(t[0].contiguous(), t[1].contiguous(), t[2].contiguous(), group_size=group_size, bits=bits)
    return ops.scaled_dot_product_attention(queries, deq(q_keys), deq(q_values), scale=scale, mask=mask)


def scaled_dot_product_attention(queries, keys, values, cache, scale, mask):
    """base.py:92-113."""
    if isinstance(cache, (QuantizedKVCache, PagedQuantizedKVCache)):
        return quantized_scaled_dot_product_attention(queries, keys, values, scale=scale, mask=mask,
                                                      group_size=cache.group_size, bits=cache.bits, cache=cache)
    if isinstance(keys, PagedKVCache):
        c = keys
        B, Hq, L, D = queries.shape
        if L > 1:   # prefill (also a continuation behind re-used prefix pages): the kernel walks the page table
            if not isinstance(mask, ops.CausalMask) or mask.L != L or mask.offset + L != c.offset:
                raise ValueError("PagedKVCache prefill needs the causal mask of create_attention_mask")
            return ops.prefill_attention(queries, c.k_pool, c.v_pool, c.layout(), mask.offset, scale)
        if mask is not None:
            raise ValueError("PagedKVCache decode step takes no mask")
        pos = torch.full((B,), c.offset - 1, dtype=torch.int32, device=queries.device)
        out = ops.decode_attention(queries.reshape(B, Hq, D).contiguous(), c.k_pool, c.v_pool, c.layout(), pos, scale,
                                   max_context=c.offset)
        return out.view(B, Hq, 1, D)
    return ops.scaled_dot_product_attention(queries, keys, values, scale=scale, mask=mask)


class Attention(nn.Module):
    def __init__(self, args: ModelArgs, device="cuda"):
        dim = args.hidden_size
        self.num_heads = args.num_attention_heads
        self.n_kv_heads = args.num_key_value_heads or args.num_attention_heads
        self.head_dim = args.head_dim or args.hidden_size // self.num_heads
        self.scale = self.head_dim ** -0.5
        q = args.quantization
        mk = lambda i, o: nn.QuantizedLinear(i, o, bias=False, group_size=q["group_size"], bits=q["bits"], device=device)
        self.q_proj = mk(dim, self.num_heads * self.head_dim)
        self.k_proj = mk(dim, self.n_kv_heads * self.head_dim)
        self.v_proj = mk(dim, self.n_kv_heads * self.head_dim)
        self.o_proj = mk(self.num_heads * self.head_dim, dim)
        max_len = args.max_position_embeddings or 8192
        rs = args.rope_scaling or {}
        self.rope = Llama3RoPE(max_len, max_len, self.head_dim, args.rope_theta, float(rs.get("factor", 1.0)),
                               float(rs.get("low_freq_factor", 1.0)), float(rs.get("high_freq_factor", 1.0)),
                               args.rope_traditional, device=device)

    def __call__(self, x, mask=None, cache=None):
        B, L, _ = x.shape
        queries, keys, values = self.q_proj(x), self.k_proj(x), self.v_proj(x)
        queries = queries.reshape(B, L, self.num_heads, -1).transpose(1, 2)
        keys = keys.reshape(B, L, self.n_kv_heads, -1).transpose(1, 2)
        values = values.reshape(B, L, self.n_kv_heads, -1).transpose(1, 2)
        if cache is not None:
            queries = self.rope(queries, offset=cache.offset)
            keys = self.rope(keys, offset=cache.offset)
            keys, values = cache.update_and_fetch(keys, values)
        else:
            queries, keys = self.rope(queries), self.rope(keys)
        output = scaled_dot_product_attention(queries, keys, values, cache=cache, scale=self.scale, mask=mask)
        output = output.transpose(1, 2).reshape(B, L, -1)
        # tensor parallel: o_proj holds a K slice; the ranks' partial sums meet in one all-reduce (tp.py)
        return all_reduce_sum(self.o_proj(output), getattr(self, "tp", None))


class MLP(nn.Module):
    def __init__(self, args: ModelArgs, device="cuda"):
        q = args.quantization
        mk = lambda i, o: nn.QuantizedLinear(i, o, bias=False, group_size=q["group_size"], bits=q["bits"], device=device)
        self.gate_proj = mk(args.hidden_size, args.intermediate_size)
        self.down_proj = mk(args.intermediate_size, args.hidden_size)
        self.up_proj = mk(args.hidden_size, args.intermediate_size)

    def __call__(self, x):
        return all_reduce_sum(self.down_proj(ops.silu_mul(self.gate_proj(x), self.up_proj(x))), getattr(self, "tp", None))


class TransformerBlock(nn.Module):
    def __init__(self, args: ModelArgs, device="cuda"):
        self.self_attn = Attention(args, device)
        self.mlp = MLP(args, device)
        self.input_layernorm = nn.RMSNorm(args.hidden_size, eps=args.rms_norm_eps, device=device)
        self.post_attention_layernorm = nn.RMSNorm(args.hidden_size, eps=args.rms_norm_eps, device=device)

    def __call__(self, x, mask=None, cache=None):
        r = self.self_attn(self.input_layernorm(x), mask, cache)
        h = ops.add(x, r)
        r = self.mlp(self.post_attention_layernorm(h))
        return ops.add(h, r)


class LlamaModel(nn.Module):
    def __init__(self, args: ModelArgs, device="cuda"):
        q = args.quantization
        self.embed_tokens = nn.QuantizedEmbedding(args.vocab_size, args.hidden_size, q["group_size"], q["bits"], device)
        self.layers = [TransformerBlock(args, device) for _ in range(args.num_hidden_layers)]
        self.norm = nn.RMSNorm(args.hidden_size, eps=args.rms_norm_eps, device=device)

    def __call__(self, inputs, mask=None, cache=None):
        h = self.embed_tokens(inputs)
        if mask is None:
            mask = create_attention_mask(h, cache)
        if cache is None:
            cache = [None] * len(self.layers)
        for layer, c in zip(self.layers, cache):
            h = layer(h, mask, cache=c)
        return self.norm(h)


class Model(nn.Module):
    def __init__(self, args: ModelArgs, device="cuda"):
        self.args = args
        self.model_type = args.model_type
        self.model = LlamaModel(args, device)
        if not args.tie_word_embeddings:
            q = args.quantization
            self.lm_head = nn.QuantizedLinear(args.hidden_size, args.lm_head_rows or args.vocab_size, bias=False,
                                              group_size=q["group_size"], bits=q["bits"], device=device)

    def set_tp(self, tp: TPInfo):
        """Mark this model as rank ``tp.rank`` of a tensor-parallel group (weights already sharded by tp.shard_weights)."""
        self.tp = tp
        for ly in self.model.layers:
            ly.self_attn.tp = tp
            ly.mlp.tp = tp
        return self

    def head_slices(self):
        """(weight, scales, biases) of the output head this rank applies: lm_head, or -- tied -- the embedding rows
        of this rank's vocabulary shard (QuantizedEmbedding.as_linear, quantized.py:109-124)."""
        if not self.args.tie_word_embeddings:
            return self.lm_head.weight, self.lm_head.scales, self.lm_head.biases
        e = self.model.embed_tokens
        tp = getattr(self, "tp", None)
        if tp is None or not tp.on:
            return e.weight, e.scales, e.biases
        n = self.args.vocab_size // tp.world
        lo, hi = tp.rank * n, (tp.rank + 1) * n
        return e.weight[lo:hi], e.scales[lo:hi], e.biases[lo:hi]

    def __call__(self, inputs, mask=None, cache=None):
        out = self.model(inputs, mask, cache)
        q = self.args.quantization
        w, s, b = self.head_slices()
        logits = ops.quantized_matmul(out, w, s, b, transpose=True, group_size=q["group_size"], bits=q["bits"])
        return all_gather_last(logits, getattr(self, "tp", None))

    def sanitize(self, weights: dict) -> dict:
        """language.py:212-219."""
        weights = {k: v for k, v in weights.items() if "self_attn.rotary_emb.inv_freq" not in k}
        if self.args.tie_word_embeddings:
            for k in [k for k in weights if k.startswith("lm_head.")]:
                weights.pop(k)
        return weights

    @property
    def layers(self):
        return self.model.layers

    def make_cache(self, capacity: int = 0, kv_bits: int | None = None, kv_group_size: int = 64):
        """One cache per layer; kv_bits=8 gives the reference's QuantizedKVCache (PIE/cache/kv_cache/quantized.py)."""
        if kv_bits is not None:
            return [QuantizedKVCache(kv_group_size, kv_bits, capacity) for _ in self.layers]
        return [KVCache(capacity) for _ in self.layers]

    def parameters(self, prefix: str = "") -> dict:
        p = super().parameters(prefix)
        return {k: v for k, v in p.items() if not k.endswith("_freqs")}
