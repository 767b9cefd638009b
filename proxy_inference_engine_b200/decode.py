"""Decode engine: the generate_step loop of PIE (PIE/engine/inference_engine.py:228-297) with temp == 0, run as
one CUDA-graph replay per token over the fused sm_100a kernels.

One step (the wiring of PIE/models/llama/language.py:75-210, L == 1), 5 launches per layer:
    h      = dequantize(embed[token])                                   QuantizedEmbedding.__call__
    per layer:
      qkv  = qmv_fused(rms_norm(h, w_in) ; [q_proj | k_proj | v_proj])   input_layernorm + 3 QuantizedLinear
      q    = rope(q), cache[pos] = (rope(k), v)                         Llama3RoPE + update_and_fetch
      o    = flash-decode(q, cache[:pos + 1])                           scaled_dot_product_attention
      h    = h + qmv(o ; o_proj)                                        o_proj + residual
      a    = silu(gate) * up,  (gate, up) = qmv(rms_norm(h, w_post))    post_attention_layernorm + MLP front
      h    = h + qmv(a ; down_proj)                                     down_proj + residual
    logits = qmv_fused(rms_norm(h, w_norm) ; lm_head)                   norm + head (tied: embed.as_linear)
    token  = argmax(logits); positions += 1; seq[b, pos] = token        sampler, cache.offset += 1
Token ids and positions stay on the device; the host reads the token list back when it wants it.
"""
from __future__ import annotations

import ctypes as C
import os

import torch
import torch.distributed as dist

from . import _lib, ops
from ._lib import EPI_RESADD, EPI_SILUMUL, EPI_STORE, KvLayout, QmvSeg, StepIO, StepLayer, StepModel
from .cache import KVCache, PagedKVCache, PagedQuantizedKVCache, QuantizedKVCache
from .llama import Model
from .ops import _p, _stream


def _segs(*linears):
    arr = (QmvSeg * len(linears))()
    for i, l in enumerate(linears):
        arr[i].wq, arr[i].scales, arr[i].biases = l.weight.data_ptr(), l.scales.data_ptr(), l.biases.data_ptr()
        arr[i].N = l.weight.shape[0]
    return arr


class DecodeEngine:
    def __init__(self, model: Model, batch: int = 1, max_context: int = 1024, paged: bool = False,
                 use_graph: bool = True, pdl: bool = True, keep_logits: bool = True, seed: int = 2,
                 fused: bool = False, trace: bool = False, attn_fused: bool | None = None, tp=None,
                 sampling: dict | None = None, kv_bits: int | None = None):
        """kv_bits=8: the KV cache is the reference's QuantizedKVCache (8 bits, groups of 64; PIE/cache/kv_cache/quantized.py)
        and the step's attention is pie_kvq_append + pie_decode_attn_q8.
        sampling: make_sampler's arguments (temp, top_p, min_p, min_tokens_to_keep, top_k; PIE/samplers/__init__.py:11)
        plus `seed`; None or temp == 0 is greedy.  With temp > 0 every step ends in pie_sample instead of pie_argmax, inside
        the captured graph (the Philox draw counter lives in device memory), and token ids still never leave the GPU."""
        a = model.args
        sp = dict(sampling or {})
        self.sampling = None
        if sp.get("temp", 0.0) > 0:
            if fused:
                raise ValueError("the one-launch decode step samples greedily; use the per-operator path with sampling")
            self.sampling = dict(temp=float(sp["temp"]), top_p=float(sp.get("top_p", 0.0)), min_p=float(sp.get("min_p", 0.0)),
                                 min_tokens_to_keep=int(sp.get("min_tokens_to_keep", 1)), top_k=int(sp.get("top_k", -1)))
            self.rng_state = ops.new_rng_state(sp.get("seed", 0), model.model.norm.weight.device)
        self.model, self.args, self.B, self.max_context = model, a, batch, max_context
        self.pdl, self.use_graph = int(pdl), use_graph
        dev = model.model.norm.weight.device
        self.dev = dev
        at0 = model.layers[0].self_attn
        self.Hq, self.Hkv, self.D = at0.num_heads, at0.n_kv_heads, at0.head_dim
        self.gs, self.bits = a.quantization["group_size"], a.quantization["bits"]
        bf = torch.bfloat16
        B, h = batch, a.hidden_size
        self.tokens = torch.zeros((B,), dtype=torch.int32, device=dev)
        self.positions = torch.zeros((B,), dtype=torch.int32, device=dev)
        self.seq = torch.zeros((B, max_context + 1), dtype=torch.int32, device=dev)
        self.h = torch.empty((B, h), dtype=bf, device=dev)
        self.qkv = torch.empty((B, (self.Hq + 2 * self.Hkv) * self.D), dtype=bf, device=dev)
        self.q = torch.empty((B, self.Hq, self.D), dtype=bf, device=dev)
        self.attn = torch.empty((B, self.Hq * self.D), dtype=bf, device=dev)
        self.act = torch.empty((B, a.intermediate_size), dtype=bf, device=dev)
        self.logits = torch.empty((B, a.vocab_size), dtype=bf, device=dev)
        self.rope_tab = torch.empty((B, 2, self.D // 2), dtype=torch.float32, device=dev)
        # tensor parallel (tp.py): this rank owns Hq/W query heads, Hkv/W kv heads, ffn/W columns and V/W logits
        self.tp = tp if tp is not None else getattr(model, "tp", None)
        self.tp_on = self.tp is not None and self.tp.on
        if self.tp_on:
            if fused:
                raise ValueError("the one-launch decode step is single-GPU; tensor parallel uses the per-operator path")
            w = self.tp.world
            self.logits_local = torch.empty((B, a.vocab_size // w), dtype=bf, device=dev)
            self.logits_gather = torch.empty((w * B, a.vocab_size // w), dtype=bf, device=dev)  # rank-major rows
            # exchange steps: one-shot pushes over NVLink peer memory (tp.TPExchange, csrc/pie_comm.cu) unless
            # PIE_TP_ONESHOT=0 or symmetric memory cannot be set up (then NCCL collectives, as in round 1)
            self.xch = None
            self.xch_f = None
            self.part = torch.empty((B, h), dtype=bf, device=dev)
            if os.environ.get("PIE_TP_ONESHOT", "1") != "0" and B * h <= 131072:
                try:
                    from .tp import TPExchange
                    self.xch = TPExchange(self.tp, B * h, B, a.vocab_size // w, dev)
                    self.logits = self.xch.gathered   # peers write their vocabulary slices straight into it
                    # PIE_TP_FUSED=1 (batch 1, 4-bit): the all-reduce FUSED into the GEMVs around it (pie_qmv_fused_tp):
                    # the o / down projection pushes its partial row to every peer, the next matrix's x load adds the
                    # slots and the residual.  Its own buffer and exchange numbers; the residual stream alternates
                    # between h and h2.  Parity-green but measured SLOWER than the stand-alone exchange kernel on 2 x B200
                    # (8B bs1: 502 vs 572 tok/s): every producer CTA pays a system-scope fence before it may signal, and
                    # that sits on the critical path of the next matrix -- hence opt-in.
                    self.xch_f = None
                    if B == 1 and self.bits == 4 and os.environ.get("PIE_TP_FUSED", "0") == "1":
                        self.xch_f = TPExchange(self.tp, h, 1, 8, dev)
                        self.h2 = torch.empty_like(self.h)
                        self.tpf_push = self.xch_f.fused(1)
                        self.tpf_to_h2 = self.xch_f.fused(2, self.h2)
                        self.tpf_to_h = self.xch_f.fused(2, self.h)
                except Exception as e:  # pragma: no cover - depends on the box
                    import warnings
                    warnings.warn(f"one-shot tensor-parallel exchange unavailable ({e!r}); using NCCL collectives")
                    self.xch = None
        # caches
        self.paged = paged
        if paged:
            # Pages come from the pool's allocator (paging.PageAllocator over csrc/pie_pages.cu: reference counts, LIFO free
            # list).  To keep the page tables as scattered as a long-running server's, the pool is first drained and handed
            # back in a seeded random order; every row then takes the pages of its whole context.
            from .paging import PageAllocator
            n_pg = (max_context + 63) // 64
            self.page_allocator = PageAllocator(B * n_pg, self.Hkv, self.D)
            g = torch.Generator(device="cpu").manual_seed(seed)
            order = torch.randperm(B * n_pg, generator=g).tolist()
            self.page_allocator.allocate_pages(B * n_pg)
            self.page_allocator.free_pages(order)
            rows = [self.page_allocator.allocate_pages(n_pg) for _ in range(B)]
            perm = torch.tensor(rows, dtype=torch.int32).to(dev)
            self.page_table = perm
        self.kv_bits = kv_bits
        if kv_bits is not None:
            if fused or tp is not None:
                raise ValueError("kv_bits: the quantised KV cache runs on the per-operator single-GPU path")
            if paged:
                self.cache = [PagedQuantizedKVCache(B * n_pg, self.Hkv, self.D, perm, dev) for _ in model.layers]
            else:
                self.cache = model.make_cache(max_context, kv_bits=kv_bits)
                for c in self.cache:
                    c.reserve(B, self.Hkv, self.D, max_context, dev)
            self.kvq = [c.desc() for c in self.cache]   # host structs read at every launch: keep them alive
            attn_fused = False
        elif paged:
            self.cache = [PagedKVCache(torch.zeros((B * n_pg, 64, self.Hkv, self.D), dtype=bf, device=dev),
                                       torch.zeros((B * n_pg, 64, self.Hkv, self.D), dtype=bf, device=dev), perm)
                          for _ in model.layers]
            self.lay = [c.layout() for c in self.cache]
            self.kbuf = [c.k_pool for c in self.cache]
            self.vbuf = [c.v_pool for c in self.cache]
        else:
            self.cache = model.make_cache(max_context)
            for c in self.cache:
                c.reserve(B, self.Hkv, self.D, max_context, dev)
            self.lay = [ops.kv_layout(c.keys) for c in self.cache]
            self.kbuf = [c.keys for c in self.cache]
            self.vbuf = [c.values for c in self.cache]
        L = _lib.load()
        self.L = L
        self.nsplit = (L.pie_decode_attn_q8_pick_splits if kv_bits is not None else L.pie_decode_attn_pick_splits)(
            B, self.Hq, self.Hkv, max_context)
        # Two forms of decode attention.  Latency form: rope + KV append + SIMT flash-decode in ONE launch
        # (pie_decode_attn_fused).  Throughput form: pie_rope_kv_append + the tensor-core kernel of pie_fa.cu, which
        # reads K/V once per kv head.  Measured per layer on B200 (tools/run_attn_compare.sh; 8B shapes, us):
        #   bs1 ctx500 10.8 / 12.9, bs1 ctx2048 16.9 / 17.4, bs1 ctx8000 40.5 / 22.2, bs4 ctx2048 32.6 / 23.1,
        #   bs32 ctx2048 203 / 77  ->  the latency form while batch x context stays within 4096 keys.
        self.attn_fused = (batch * max_context <= 4096) if attn_fused is None else attn_fused
        # the splits of a unit form one thread-block cluster and merge through distributed shared memory; measured on
        # B200 (8B, bs1): 3..6 splits cost the same, clusters of 8 are markedly slower to schedule -> cap at 6
        self.nsplit_fused = min(6, L.pie_decode_step_pick_splits(B, self.Hq, self.Hkv, max_context))
        if os.environ.get("PIE_ATTN_SPLITS"):  # tuning knob (tools/bench_step_parts.py)
            self.nsplit_fused = int(os.environ["PIE_ATTN_SPLITS"])
        nb = L.pie_decode_attn_workspace_bytes(B, self.Hq, self.D, max(self.nsplit, self.nsplit_fused))
        self.attn_ws = torch.zeros(max(nb, 256), dtype=torch.uint8, device=dev)
        if os.environ.get("PIE_COLOCATE", "1") != "0":
            from .checkpoint import colocate_for_decode
            colocate_for_decode(model)   # fewer TMA copies per weight tile (see checkpoint.colocate_for_decode)
        # per-layer segment tables (host structs must outlive the captured graph: keep references)
        self.seg_qkv, self.seg_o, self.seg_gu, self.seg_down = [], [], [], []
        nb = 0
        mb = min(B, 32)
        for ly in model.layers:
            at, mlp = ly.self_attn, ly.mlp
            self.seg_qkv.append(_segs(at.q_proj, at.k_proj, at.v_proj))
            self.seg_o.append(_segs(at.o_proj))
            self.seg_gu.append(_segs(mlp.gate_proj, mlp.up_proj))
            self.seg_down.append(_segs(mlp.down_proj))
        hw, hs, hb = model.head_slices()
        self._head_keep = (hw, hs, hb)
        self.seg_head = (QmvSeg * 1)()
        self.seg_head[0].wq, self.seg_head[0].scales, self.seg_head[0].biases = hw.data_ptr(), hs.data_ptr(), hb.data_ptr()
        self.seg_head[0].N = hw.shape[0]
        qd = self.Hq * self.D
        for (n, k, dual) in (((self.Hq + 2 * self.Hkv) * self.D, h, 0), (h, qd, 0), (a.intermediate_size, h, 1),
                             (h, a.intermediate_size, 0), (hw.shape[0], h, 0)):
            nb = max(nb, L.pie_qmv_workspace_bytes(mb, n, k, self.bits, dual, 0))
        self.qmv_ws = torch.zeros(max(nb, 256), dtype=torch.uint8, device=dev)
        self.graph = None
        self._next_pos = None
        self.launches_per_step = 0
        # fused=True: one cooperative launch per step (pie_decode_step), B <= 8.  Round-1 status: parity-green but
        # slower than the per-operator launches, hence opt-in
        self.fused = bool(fused)
        self.trace = trace
        if self.fused:
            self._plan_fused()

    def _plan_fused(self):
        """Describe the step to pie_decode_step_plan (host only) and upload the plan once."""
        a, m, L, B = self.args, self.model, self.L, self.B
        if B != 1 or self.bits != 4:
            raise ValueError("the one-launch decode step serves batch 1 with 4-bit weights (the latency-bound case)")

        def seg(l):
            return QmvSeg(l.weight.data_ptr(), l.scales.data_ptr(), l.biases.data_ptr(), l.weight.shape[0])

        layers = (StepLayer * len(m.layers))()
        for i, ly in enumerate(m.layers):
            at, mlp, s = ly.self_attn, ly.mlp, layers[i]
            s.q, s.k, s.v, s.o = seg(at.q_proj), seg(at.k_proj), seg(at.v_proj), seg(at.o_proj)
            s.gate, s.up, s.down = seg(mlp.gate_proj), seg(mlp.up_proj), seg(mlp.down_proj)
            s.input_norm = ly.input_layernorm.weight.data_ptr()
            s.post_norm = ly.post_attention_layernorm.weight.data_ptr()
            s.kcache, s.vcache, s.kv = self.kbuf[i].data_ptr(), self.vbuf[i].data_ptr(), self.lay[i]
        at0 = m.layers[0].self_attn
        head = m.model.embed_tokens if a.tie_word_embeddings else m.lm_head
        sm = StepModel(len(m.layers), a.hidden_size, a.intermediate_size, self.Hq, self.Hkv, self.D, a.vocab_size,
                       self.bits, self.gs, float(a.rms_norm_eps), float(at0.scale), at0.rope._freqs.data_ptr(),
                       seg(m.model.embed_tokens), seg(head), m.model.norm.weight.data_ptr())
        self.step_nsplit = L.pie_decode_step_pick_splits(B, self.Hq, self.Hkv, self.max_context)
        nb = L.pie_decode_step_attn_ws_bytes(B, self.Hq, self.D, self.step_nsplit)
        self.step_attn_ws = torch.zeros(max(nb, 256), dtype=torch.uint8, device=self.dev)
        self.step_sync = torch.zeros(256, dtype=torch.uint8, device=self.dev)
        self.step_trace = (torch.zeros(4 + 7 * (4 * len(m.layers) + 1), dtype=torch.int64, device=self.dev)
                           if self.trace else None)
        io = StepIO(B, _p(self.tokens), _p(self.positions), _p(self.seq), self.seq.shape[1], _p(self.h), _p(self.qkv),
                    _p(self.attn), _p(self.act), _p(self.logits), _p(self.step_attn_ws), self.step_attn_ws.numel(),
                    self.step_nsplit, _p(self.qmv_ws), self.qmv_ws.numel(), _p(self.step_trace))
        nbytes = L.pie_decode_step_plan_bytes(len(m.layers))
        self.plan_host = (C.c_uint8 * nbytes)()
        _lib.check(L.pie_decode_step_plan(C.byref(sm), layers, C.byref(io), self.plan_host, nbytes), "decode_step_plan")
        self.plan_dev = torch.frombuffer(self.plan_host, dtype=torch.uint8).clone().to(self.dev)
        self._plan_keep = (layers, sm, io)

    # ------------------------------------------------------------------------------------------------
    def _qmv(self, out, out_ld, x, x_ld, K, segs, nseg, norm_w, epi, residual, tpf=None):
        _lib.check(self.L.pie_qmv_fused_tp(_p(out), out_ld, _p(x), x_ld, self.B, K, segs, nseg, self.gs, self.bits,
                                           _p(norm_w), float(self.args.rms_norm_eps), epi, _p(residual), 0,
                                           _p(self.qmv_ws), self.qmv_ws.numel(), self.pdl,
                                           C.byref(tpf) if tpf is not None else None, _stream()), "qmv_fused")

    def _qmv_rowsplit(self, h, hid, x, K, segs):
        """h += x @ W^T for a matrix split by input columns (o_proj, down_proj).  Single GPU: residual-add epilogue.
        Tensor parallel: rank 0 folds the residual into its partial sum, the others store theirs, and ONE all-reduce
        over [B, hidden] leaves the new residual stream on every rank (the exchange step of
        QuantizedShardedToAllLinear, MLX/python/mlx/nn/layers/distributed.py:459-566)."""
        if self.tp_on and self.xch is not None:
            # every rank stores its bare partial sums; the exchange adds them in rank order and THEN the residual
            # (the order of MLX's sharded linear + TransformerBlock: all_sum first, h + r after)
            self._qmv(self.part, hid, x, K, K, segs, 1, None, EPI_STORE, None)
            self.xch.all_reduce(h, self.part, h, _stream())
            return
        if not self.tp_on or self.tp.rank == 0:
            self._qmv(h, hid, x, K, K, segs, 1, None, EPI_RESADD, h)
        else:
            self._qmv(h, hid, x, K, K, segs, 1, None, EPI_STORE, None)
        if self.tp_on:
            dist.all_reduce(h, op=dist.ReduceOp.SUM, group=self.tp.group)

    def _step_kernels(self):
        """Enqueue one decode step on the current stream (capturable: no sync, no allocation)."""
        a, m, L = self.args, self.model, self.L
        if self.fused:
            _lib.check(L.pie_decode_step(_p(self.plan_dev), self.plan_host, len(self.plan_host), _p(self.step_sync),
                                         _stream()), "decode_step")
            return
        emb = m.model.embed_tokens
        h, hid = self.h, a.hidden_size
        _lib.check(L.pie_dequantize(_p(h), _p(emb.weight), _p(emb.scales), _p(emb.biases), _p(self.tokens), self.B,
                                    emb.weight.shape[0], hid, self.gs, self.bits, _stream()), "embed")
        if self.attn_fused:   # this step's rotary sines, once for all layers
            _lib.check(L.pie_rope_table(_p(self.rope_tab), _p(self.positions), _p(m.layers[0].self_attn.rope._freqs),
                                        self.B, self.D // 2, _stream()), "rope_table")
        qd = self.Hq * self.D
        import os
        skip = set(filter(None, os.environ.get("PIE_DEBUG_SKIP", "").split(",")))  # tools/bench_step_parts.py only
        fz = self.tp_on and getattr(self, "xch_f", None) is not None   # all-reduces fused into the GEMVs (batch 1)
        for i, ly in enumerate(m.layers):
            if "qkv" not in skip:
                if fz and i > 0:   # x = h2 + (down partials of the previous layer); the new stream lands in h
                    self._qmv(self.qkv, self.qkv.shape[1], self.h2, hid, hid, self.seg_qkv[i], 3,
                              ly.input_layernorm.weight, EPI_STORE, None, self.tpf_to_h)
                else:
                    self._qmv(self.qkv, self.qkv.shape[1], h, hid, hid, self.seg_qkv[i], 3, ly.input_layernorm.weight,
                              EPI_STORE, None)
            if "attn" not in skip and self.attn_fused:
                _lib.check(L.pie_decode_attn_fused(_p(self.attn), _p(self.qkv), _p(self.kbuf[i]), _p(self.vbuf[i]),
                                                   C.byref(self.lay[i]), _p(self.positions),
                                                   _p(ly.self_attn.rope._freqs), _p(self.rope_tab), self.B, self.Hq,
                                                   self.Hkv, self.D,
                                                   float(ly.self_attn.scale), self.nsplit_fused, _p(self.attn_ws),
                                                   self.attn_ws.numel(), self.pdl, _stream()), "decode_attn_fused")
            if self.kv_bits is not None:
                if "rope" not in skip:
                    _lib.check(L.pie_kvq_append(_p(self.q), _p(self.qkv), C.byref(self.kvq[i]), _p(self.positions),
                                                _p(ly.self_attn.rope._freqs), self.B, self.Hq, self.Hkv, self.D,
                                                _stream()), "kvq_append")
                if "attn" not in skip:
                    _lib.check(L.pie_decode_attn_q8(_p(self.attn), _p(self.q), C.byref(self.kvq[i]), _p(self.positions),
                                                    self.B, self.Hq, self.Hkv, self.D, float(ly.self_attn.scale),
                                                    self.nsplit, _p(self.attn_ws), self.attn_ws.numel(), _stream()),
                               "decode_attn_q8")
            elif "rope" not in skip and not self.attn_fused:
                _lib.check(L.pie_rope_kv_append(_p(self.q), _p(self.qkv), _p(self.kbuf[i]), _p(self.vbuf[i]),
                                            C.byref(self.lay[i]), _p(self.positions), _p(ly.self_attn.rope._freqs),
                                                self.B, self.Hq, self.Hkv, self.D, _stream()), "rope_kv_append")
            if "attn" not in skip and not self.attn_fused and self.kv_bits is None:
                _lib.check(L.pie_decode_attn(_p(self.attn), _p(self.q), _p(self.kbuf[i]), _p(self.vbuf[i]),
                                         C.byref(self.lay[i]), _p(self.positions), self.B, self.Hq, self.Hkv, self.D,
                                         float(ly.self_attn.scale), self.nsplit, _p(self.attn_ws),
                                         self.attn_ws.numel(), _stream()), "decode_attn")
            if fz:
                self._qmv(self.part, hid, self.attn, qd, qd, self.seg_o[i], 1, None, EPI_STORE, None, self.tpf_push)
                self._qmv(self.act, a.intermediate_size, h, hid, hid, self.seg_gu[i], 2,
                          ly.post_attention_layernorm.weight, EPI_SILUMUL, None, self.tpf_to_h2)
                self._qmv(self.part, hid, self.act, a.intermediate_size, a.intermediate_size, self.seg_down[i], 1, None,
                          EPI_STORE, None, self.tpf_push)
                continue
            if "o" not in skip:
                self._qmv_rowsplit(h, hid, self.attn, qd, self.seg_o[i])
            if "gu" not in skip:
                self._qmv(self.act, a.intermediate_size, h, hid, hid, self.seg_gu[i], 2,
                          ly.post_attention_layernorm.weight, EPI_SILUMUL, None)
            if "down" not in skip:
                self._qmv_rowsplit(h, hid, self.act, a.intermediate_size, self.seg_down[i])
        if self.tp_on:
            nl = self.logits_local.shape[1]
            if fz:   # x = h2 + (down partials of the last layer)
                self._qmv(self.logits_local, nl, self.h2, hid, hid, self.seg_head, 1, m.model.norm.weight, EPI_STORE, None,
                          self.tpf_to_h)
            else:
                self._qmv(self.logits_local, nl, h, hid, hid, self.seg_head, 1, m.model.norm.weight, EPI_STORE, None)
            if self.xch is not None:
                self.xch.all_gather(self.logits_local, _stream())
                self._sample()
                return
            dist.all_gather_into_tensor(self.logits_gather, self.logits_local, group=self.tp.group)
            self.logits.view(self.B, self.tp.world, nl).copy_(
                self.logits_gather.view(self.tp.world, self.B, nl).permute(1, 0, 2))
        else:
            self._qmv(self.logits, a.vocab_size, h, hid, hid, self.seg_head, 1, m.model.norm.weight, EPI_STORE, None)
        self._sample()

    def _sample(self, logits=None):
        """tokens = sampler(logits); positions += 1; seq[b, pos] = token   (inference_engine.py:268-271), one launch."""
        lg = self.logits if logits is None else logits
        if self.sampling is None:
            ops.argmax(lg, positions_to_advance=self.positions, seq_out=self.seq, out=self.tokens)
        else:
            ops.sample(lg, rng_state=self.rng_state, positions_to_advance=self.positions, seq_out=self.seq,
                       out=self.tokens, **self.sampling)

    def enqueue_gemvs(self):
        """Only the quantized matmuls of one step (4 per layer + head), same descriptors, no collectives: bench.py times
        a CUDA graph of these to get the dominant kernel's in-graph launch duration (the outputs are scratch)."""
        a, m = self.args, self.model
        h, hid, qd = self.h, a.hidden_size, self.Hq * self.D
        for i, ly in enumerate(m.layers):
            self._qmv(self.qkv, self.qkv.shape[1], h, hid, hid, self.seg_qkv[i], 3, ly.input_layernorm.weight, EPI_STORE, None)
            self._qmv(h, hid, self.attn, qd, qd, self.seg_o[i], 1, None, EPI_RESADD, h)
            self._qmv(self.act, a.intermediate_size, h, hid, hid, self.seg_gu[i], 2, ly.post_attention_layernorm.weight,
                      EPI_SILUMUL, None)
            self._qmv(h, hid, self.act, a.intermediate_size, a.intermediate_size, self.seg_down[i], 1, None, EPI_RESADD, h)
        out = self.logits_local if self.tp_on else self.logits
        self._qmv(out, out.shape[1], h, hid, hid, self.seg_head, 1, m.model.norm.weight, EPI_STORE, None)
        return 4 * len(m.layers) + 1

    def enqueue_exchanges(self):
        """Only the tensor-parallel exchange steps of one decode step (2 per layer + the logits gather): bench.py times a
        CUDA graph of these (the data exchanged is scratch).  Returns how many."""
        n = 0
        for _ in self.model.layers:
            for _ in range(2):
                if self.xch is not None:
                    self.xch.all_reduce(self.part, self.part, None, _stream())
                else:
                    dist.all_reduce(self.part, op=dist.ReduceOp.SUM, group=self.tp.group)
                n += 1
        if self.xch is not None:
            self.xch.all_gather(self.logits_local, _stream())
        else:
            dist.all_gather_into_tensor(self.logits_gather, self.logits_local, group=self.tp.group)
        return n + 1

    def _ensure_graph(self):
        if not self.use_graph or self.graph is not None:
            return
        # warm-up (also opts the kernels into their shared-memory size) on a side stream, then capture.
        # The step mutates tokens / positions / cache: save and restore them around warm-up and capture.
        saved = (self.tokens.clone(), self.positions.clone(), self.seq.clone())
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            n0 = ops.launch_count()
            self._step_kernels()
            self.launches_per_step = ops.launch_count() - n0
            self._restore(saved)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._step_kernels()
        self._restore(saved)
        self.graph = g

    def _restore(self, saved):
        self.tokens.copy_(saved[0]); self.positions.copy_(saved[1]); self.seq.copy_(saved[2])

    # ------------------------------------------------------------------------------------------------
    def prefill(self, prompt_ids: torch.Tensor):
        """Whole prompt in ONE forward (inference_engine.py:278), B rows of equal length; samples the first token."""
        ids = prompt_ids.to(self.dev).to(torch.int32).reshape(self.B, -1)
        Lp = ids.shape[1]
        if Lp + 1 > self.max_context:
            raise ValueError("prompt longer than the engine's max_context")
        if self.paged and self.kv_bits is None:
            tmp = self.model.make_cache(Lp)
            logits = self.model(ids, cache=tmp)
            for c, t in zip(self.cache, tmp):
                c.offset = 0
                c.update_and_fetch(t.keys[:, :, :Lp], t.values[:, :, :Lp])
        else:
            for c in self.cache:
                c.offset = 0
            logits = self.model(ids, cache=self.cache)
        self.positions.fill_(Lp - 1)
        self._next_pos = Lp  # host-side mirror of the device positions: the cache slot the next step writes
        self.seq.zero_()
        self.seq[:, :Lp] = ids
        self.prompt_len = Lp
        last = logits[:, -1, :].contiguous()
        self.logits.copy_(last)
        self._sample(self.logits)
        return self.tokens

    def step(self, n: int = 1):
        """n decode steps; each consumes self.tokens at self.positions and samples the next token.
        The KV buffers hold max_context tokens: a step past that would write another row's (or page's) K/V, so it is
        refused here (the reference grows its cache instead, reusable.py:167-203; this engine's buffers are fixed
        because the captured graph holds their addresses)."""
        if getattr(self, "_next_pos", None) is None:
            raise ValueError("DecodeEngine.step() before prefill()")
        if self._next_pos + n > self.max_context:
            raise ValueError(f"decode step past max_context: position {self._next_pos} + {n} steps > {self.max_context}")
        self._next_pos += n
        self._ensure_graph()
        for _ in range(n):
            if self.graph is not None:
                self.graph.replay()
            else:
                n0 = ops.launch_count()
                self._step_kernels()
                self.launches_per_step = ops.launch_count() - n0
        for c in self.cache:
            c.offset += n

    def generated(self) -> torch.Tensor:
        """Tokens sampled so far, [B, n] (device -> host copy: the sync point, inference_engine.py:202)."""
        end = int(self.positions[0].item()) + 1
        return self.seq[:, self.prompt_len:end].cpu()

    def generate(self, prompt_ids, n_steps: int) -> torch.Tensor:
        self.prefill(prompt_ids)
        self.step(n_steps - 1)
        return self.generated()
