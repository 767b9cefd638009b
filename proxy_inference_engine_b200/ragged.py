"""Ragged continuous batching: one model step over a mix of prefill chunks and decode rows.

``BatchDetails`` carries the fields of the reference's descriptor (CORE/include/engine/batch_details.hpp:10-90): the
concatenated ``token_ids`` / ``positions`` of the step, per-sequence ``input_lengths`` / ``context_lengths``, the
``consolidated_block_table`` of page ids (this build: 2-D ``[num_sequences, max_blocks]``, pages of 64 tokens from
``paging.PageAllocator``) and the prefill / decode counts.  The reference declares the descriptor but no kernel consumes
it yet (its scheduler is a skeleton); here ``ragged_step`` runs it:

* every token of the step is one row of the activation matrix, so the quantized matmuls see M = total_tokens_in_step;
* ``pie_rope_kv_append`` rotates q and k at each token's own position and appends k, v through the token's row of a
  per-token block table (the sequence's row, repeated) -- per-sequence positions, no uniform ``offset``
  (lifting PIE/cache/kv_cache/reusable.py:139);
* ``pie_decode_attn`` then gives every token the context ``position + 1`` of its own sequence: inside a prefill chunk
  that is exactly the causal mask of create_attention_mask (PIE/models/base.py:18-53), because the chunk's keys were
  appended before any of its queries attends.

Same arithmetic as the per-sequence path (Model.__call__ over a PagedKVCache); tests/test_gpu_ragged.py compares the two.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import torch

from . import _lib, ops
from .ops import _p, _stream

PAGE_TOKENS = 64


@dataclass
class BatchDetails:
    token_ids: torch.Tensor                     # int32 [total_tokens_in_step]
    positions: torch.Tensor                     # int32 [total_tokens_in_step]
    sequence_ids: list = field(default_factory=list)
    input_lengths: list = field(default_factory=list)
    context_lengths: list = field(default_factory=list)
    consolidated_block_table: torch.Tensor | None = None   # int32 [num_sequences, max_blocks]
    num_prefill_sequences: int = 0
    num_decode_sequences: int = 0
    total_tokens_in_step: int = 0

    @classmethod
    def build(cls, sequences, device="cuda") -> "BatchDetails":
        """sequences: iterable of (sequence_id, new_token_ids, context_length, page_ids).  ``context_length`` = tokens of
        the sequence already in the cache; ``page_ids`` must cover context_length + len(new_token_ids) tokens."""
        ids, pos, sids, ins, ctxs, rows = [], [], [], [], [], []
        n_pre = n_dec = 0
        for sid, toks, ctx, pages in sequences:
            toks = [int(t) for t in toks]
            if not toks:
                raise ValueError(f"sequence {sid}: no tokens in this step")
            if ctx < 0:
                raise ValueError(f"sequence {sid}: negative context length")
            if sid in sids:
                raise ValueError(f"sequence {sid} appears twice in one step")
            need = (ctx + len(toks) + PAGE_TOKENS - 1) // PAGE_TOKENS
            if len(pages) < need:
                raise ValueError(f"sequence {sid}: {len(pages)} pages for {ctx + len(toks)} tokens (needs {need})")
            ids += toks
            pos += list(range(ctx, ctx + len(toks)))
            sids.append(sid); ins.append(len(toks)); ctxs.append(int(ctx)); rows.append([int(p) for p in pages])
            if len(toks) > 1 or ctx == 0:
                n_pre += 1
            else:
                n_dec += 1
        if not sids:
            raise ValueError("empty batch")
        width = max(len(r) for r in rows)
        table = torch.tensor([r + [0] * (width - len(r)) for r in rows], dtype=torch.int32)
        return cls(torch.tensor(ids, dtype=torch.int32).to(device), torch.tensor(pos, dtype=torch.int32).to(device), sids,
                   ins, ctxs, table.to(device), n_pre, n_dec, len(ids))

    def validate(self) -> None:
        n = len(self.sequence_ids)
        if len(self.input_lengths) != n or len(self.context_lengths) != n:
            raise ValueError("BatchDetails: per-sequence vectors differ in length")
        if sum(self.input_lengths) != self.total_tokens_in_step or self.token_ids.numel() != self.total_tokens_in_step \
                or self.positions.numel() != self.total_tokens_in_step:
            raise ValueError("BatchDetails: total_tokens_in_step != sum(input_lengths) / len(token_ids)")
        if self.num_prefill_sequences + self.num_decode_sequences != n:
            raise ValueError("BatchDetails: prefill + decode counts != number of sequences")
        if self.consolidated_block_table is None or self.consolidated_block_table.shape[0] != n:
            raise ValueError("BatchDetails: block table needs one row per sequence")
        width = self.consolidated_block_table.shape[1]
        for sid, inp, ctx in zip(self.sequence_ids, self.input_lengths, self.context_lengths):
            if inp <= 0 or ctx < 0 or (ctx + inp + PAGE_TOKENS - 1) // PAGE_TOKENS > width:
                raise ValueError(f"BatchDetails: sequence {sid} does not fit its block-table row")

    def last_token_rows(self) -> torch.Tensor:
        """Row of each sequence's last token in the step (the rows whose logits the sampler reads)."""
        ends, acc = [], 0
        for n in self.input_lengths:
            acc += n
            ends.append(acc - 1)
        return torch.tensor(ends, dtype=torch.long, device=self.token_ids.device)

    def token_block_table(self) -> torch.Tensor:
        """[total_tokens, max_blocks]: the block-table row of each token's sequence (what the kernels index by row)."""
        seq_of_token = torch.repeat_interleave(
            torch.arange(len(self.input_lengths), device=self.token_ids.device),
            torch.tensor(self.input_lengths, device=self.token_ids.device))
        return self.consolidated_block_table[seq_of_token].contiguous()


def ragged_step(model, batch: BatchDetails, pools, all_logits: bool = False) -> torch.Tensor:
    """One forward step of ``model`` over ``batch``.  pools: per layer (k_pool, v_pool) bf16 [n_pages, 64, n_kv, D] --
    the pools the block table's page ids index -- or, for the 8-bit KV cache of cache.PagedQuantizedKVCache, per layer
    ((k_quants, k_scales, k_biases), (v_quants, v_scales, v_biases)) [n_pages, n_kv, 64, .].
    Returns logits [num_sequences, vocab] of each sequence's last token (all_logits: [total_tokens, vocab])."""
    batch.validate()
    L = _lib.load()
    a = model.args
    T = batch.total_tokens_in_step
    at0 = model.layers[0].self_attn
    Hq, Hkv, D = at0.num_heads, at0.n_kv_heads, at0.head_dim
    table = batch.token_block_table()
    pos = batch.positions
    h = model.model.embed_tokens(batch.token_ids[None])          # [1, T, hidden]
    q = torch.empty((T, Hq, D), dtype=torch.bfloat16, device=h.device)
    max_ctx = max(c + n for c, n in zip(batch.context_lengths, batch.input_lengths))
    for ly, (kp, vp) in zip(model.layers, pools):
        at = ly.self_attn
        x = ly.input_layernorm(h)
        qkv = torch.cat([at.q_proj(x), at.k_proj(x), at.v_proj(x)], dim=-1).view(T, (Hq + 2 * Hkv) * D)
        if isinstance(kp, (tuple, list)):   # quantised pools: (quants, scales, biases) [n_pages, n_kv, 64, .] per K and V
            desc = ops.kvq_desc(tuple(kp), tuple(vp), table, PAGE_TOKENS)
            ops.kvq_append(qkv, desc, pos, at.rope._freqs, Hq, Hkv, D, q_out=q)
            o = ops.decode_attention_q8(q, desc, pos, at.scale, Hkv, max_context=max_ctx)
        else:
            lay = ops.kv_layout(kp, table, PAGE_TOKENS)
            _lib.check(L.pie_rope_kv_append(_p(q), _p(qkv), _p(kp), _p(vp), C.byref(lay), _p(pos), _p(at.rope._freqs), T,
                                            Hq, Hkv, D, _stream()), "rope_kv_append")
            o = ops.decode_attention(q, kp, vp, lay, pos, at.scale, max_context=max_ctx)     # [T, Hq * D]
        h = ops.add(h, at.o_proj(o[None]))
        h = ops.add(h, ly.mlp(ly.post_attention_layernorm(h)))
    if not all_logits:
        h = h[:, batch.last_token_rows()]
    out = model.model.norm(h.contiguous())
    w, s, b = model.head_slices()
    qz = a.quantization
    return ops.quantized_matmul(out, w, s, b, transpose=True, group_size=qz["group_size"], bits=qz["bits"])[0]
