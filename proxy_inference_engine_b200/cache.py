"""KV caches of the path.

``KVCache`` keeps the semantics of PIE's default cache (ReusableKVCache.update_and_fetch,
PIE/cache/kv_cache/reusable.py:96-142; KVCache, cache.py:28-68): buffers ``[B, n_kv, cap, D]`` growing in steps
of 256 tokens, a scalar ``offset``, ``update_and_fetch(k, v)`` returning views of the first ``offset`` tokens.
``PagedKVCache`` is the layout BASELINE config 3 names (64-token pages as CORE/include/engine/page.hpp:14,29-30,
bf16): a page pool ``[n_pages, 64, n_kv, D]`` per layer plus an int32 page table ``[B, max_pages]``.
"""
from __future__ import annotations

import torch

from . import _lib, ops
from .ops import _p, _stream
import ctypes as C


class KVCache:
    step = 256

    def __init__(self, capacity: int = 0):
        self.keys = None
        self.values = None
        self.offset = 0
        self._reserve = capacity

    def _grow(self, B, H, D, needed, device):
        cap = max(self._reserve, self.offset + needed)
        cap = ((cap + self.step - 1) // self.step) * self.step
        nk = torch.zeros((B, H, cap, D), dtype=torch.bfloat16, device=device)
        nv = torch.zeros_like(nk)
        if self.keys is not None:
            nk[:, :, : self.offset] = self.keys[:, :, : self.offset]
            nv[:, :, : self.offset] = self.values[:, :, : self.offset]
        self.keys, self.values = nk, nv

    def reserve(self, B, H, D, capacity, device="cuda"):
        self._reserve = capacity
        if self.keys is None or self.keys.shape[2] < capacity:
            self._grow(B, H, D, 0, device)

    def update_and_fetch(self, keys, values):
        B, H, L, D = keys.shape
        if self.keys is None or self.offset + L > self.keys.shape[2]:
            self._grow(B, H, D, L, keys.device)
        lay = ops.kv_layout(self.keys)
        kc, vc = keys.contiguous(), values.contiguous()  # keep both alive across the launch (no pointer reuse)
        _lib.check(_lib.load().pie_kv_write(_p(self.keys), _p(self.values), C.byref(lay), _p(kc), _p(vc), B, H, L, D,
                                            self.offset, _stream()), "update_and_fetch")
        self.offset += L
        return self.keys[:, :, : self.offset], self.values[:, :, : self.offset]

    @property
    def state(self):
        return self.keys[:, :, : self.offset], self.values[:, :, : self.offset]

    def trim(self, n: int) -> int:
        n = min(self.offset, n)
        self.offset -= n
        return n

    def is_trimmable(self) -> bool:
        return True


class PagedKVCache:
    """bf16 paged cache for one layer.  Logical token t of row b lives in page ``page_table[b, t // 64]`` at slot
    ``t % 64``.  ``update_and_fetch`` has the reference's uniform-offset contract (reusable.py:139)."""

    page_size = 64

    def __init__(self, k_pool: torch.Tensor, v_pool: torch.Tensor, page_table: torch.Tensor):
        assert k_pool.dim() == 4 and k_pool.shape[1] == self.page_size
        self.k_pool, self.v_pool, self.page_table = k_pool, v_pool, page_table
        self.offset = 0

    def layout(self):
        return ops.kv_layout(self.k_pool, self.page_table, self.page_size)

    def update_and_fetch(self, keys, values):
        B, H, L, D = keys.shape
        if self.offset + L > self.page_table.shape[1] * self.page_size:
            raise ValueError("PagedKVCache: page table too small for this many tokens")
        lay = self.layout()
        kc, vc = keys.contiguous(), values.contiguous()  # keep both alive across the launch (no pointer reuse)
        _lib.check(_lib.load().pie_kv_write(_p(self.k_pool), _p(self.v_pool), C.byref(lay), _p(kc), _p(vc), B, H, L, D,
                                            self.offset, _stream()), "update_and_fetch")
        self.offset += L
        return self, self

    def gather(self):
        """Logical [B, n_kv, offset, D] copies (tests only: the kernels read the pages in place)."""
        B = self.page_table.shape[0]
        t = torch.arange(self.offset, device=self.k_pool.device)
        pages = self.page_table[:, t // self.page_size].long()  # [B, T]
        slot = (t % self.page_size)[None].expand(B, -1)
        k = self.k_pool[pages, slot]  # [B, T, H, D]
        v = self.v_pool[pages, slot]
        return k.permute(0, 2, 1, 3).contiguous(), v.permute(0, 2, 1, 3).contiguous()


class QuantizedKVCache:
    """QuantizedKVCache (PIE/cache/kv_cache/quantized.py:8-140): keys / values are (quants uint32 [B, n_kv, cap, D*bits/32],
    scales, biases bf16 [B, n_kv, cap, D/group_size]) triples growing in steps of 256 tokens; ``update_and_fetch``
    quantises the new rows with mx.quantize's arithmetic (pie_kvq_write, bit-exact) and returns the triples cut to
    ``offset``.  8 bits, groups of 64 are built here; anything else raises as an unsupported configuration."""
    step = 256

    def __init__(self, group_size: int = 64, bits: int = 8, capacity: int = 0):
        if group_size != 64 or bits != 8:
            raise ValueError(f"QuantizedKVCache: group_size={group_size}, bits={bits} not built here (64, 8)")
        self.keys = None
        self.values = None
        self.offset = 0
        self.group_size = group_size
        self.bits = bits
        self._reserve = capacity

    def _grow(self, B, H, D, needed, device):
        cap = max(self._reserve, self.offset + needed)
        cap = ((cap + self.step - 1) // self.step) * self.step
        el = 32 // self.bits

        def fresh():
            return (torch.zeros((B, H, cap, D // el), dtype=torch.int32, device=device),
                    torch.zeros((B, H, cap, D // self.group_size), dtype=torch.bfloat16, device=device),
                    torch.zeros((B, H, cap, D // self.group_size), dtype=torch.bfloat16, device=device))
        nk, nv = fresh(), fresh()
        if self.keys is not None:
            for new, old in zip(nk + nv, self.keys + self.values):
                new[:, :, : self.offset] = old[:, :, : self.offset]
        self.keys, self.values = nk, nv

    def reserve(self, B, H, D, capacity, device="cuda"):
        self._reserve = capacity
        if self.keys is None or self.keys[0].shape[2] < capacity:
            self._grow(B, H, D, 0, device)

    def desc(self):
        return ops.kvq_desc(self.keys, self.values, group_size=self.group_size, bits=self.bits)

    def update_and_fetch(self, keys, values):
        B, H, L, D = keys.shape
        if self.keys is None or self.offset + L > self.keys[0].shape[2]:
            self._grow(B, H, D, L, keys.device)
        ops.kvq_write(self.desc(), keys, values, self.offset)
        self.offset += L
        return self.state

    @property
    def state(self):
        if self.keys is None:
            return None, None
        cut = lambda t: t[:, :, : self.offset]
        return tuple(map(cut, self.keys)), tuple(map(cut, self.values))

    def trim(self, n: int) -> int:
        n = min(self.offset, n)
        self.offset -= n
        return n

    def is_trimmable(self) -> bool:
        return True


class PagedQuantizedKVCache:
    """Quantised paged cache for one layer: pools [n_pages, n_kv, 64, D/4 | D/64] (head-major inside a page, so the 64
    tokens of a (page, head) are one contiguous 8 KB run of quants) + an int32 page table [B, max_pages]."""
    page_size = 64
    group_size, bits = 64, 8

    def __init__(self, n_pages: int, n_kv: int, D: int, page_table: torch.Tensor, device="cuda"):
        def pool():
            return (torch.zeros((n_pages, n_kv, self.page_size, D // 4), dtype=torch.int32, device=device),
                    torch.zeros((n_pages, n_kv, self.page_size, D // 64), dtype=torch.bfloat16, device=device),
                    torch.zeros((n_pages, n_kv, self.page_size, D // 64), dtype=torch.bfloat16, device=device))
        self.keys, self.values = pool(), pool()
        self.page_table = page_table
        self.offset = 0

    def desc(self):
        return ops.kvq_desc(self.keys, self.values, self.page_table, self.page_size)

    def update_and_fetch(self, keys, values):
        L = keys.shape[2]
        if self.offset + L > self.page_table.shape[1] * self.page_size:
            raise ValueError("PagedQuantizedKVCache: page table too small for this many tokens")
        ops.kvq_write(self.desc(), keys, values, self.offset)
        self.offset += L
        return self, self

    def gather(self):
        """Logical (quants, scales, biases) [B, n_kv, offset, .] copies of K and V (tests and the L > 1 attention)."""
        t = torch.arange(self.offset, device=self.page_table.device)
        pages = self.page_table[:, t // self.page_size].long()   # [B, T]
        slot = (t % self.page_size)[None].expand(pages.shape[0], -1)
        pick = lambda pool: pool[pages, :, slot].permute(0, 2, 1, 3).contiguous()   # [B, T, H, .] -> [B, H, T, .]
        return tuple(map(pick, self.keys)), tuple(map(pick, self.values))
