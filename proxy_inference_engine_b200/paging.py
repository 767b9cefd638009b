"""Paged KV cache management: the page allocator and prefix reuse.

    PageAllocator      -- CORE/include/engine/page_allocator.hpp:17-72 (same method names and failure behaviour) over the
                          lock-free C++ allocator of csrc/pie_pages.cu.
    PagedPromptCache   -- PromptCache.reuse_cache (PIE/cache/prompt_cache.py:52-76) + ReusableKVCache.reuse
                          (PIE/cache/kv_cache/reusable.py:44-94) for the paged cache: a new prompt that shares a prefix
                          with the cached one SHARES the prefix's full pages (reference count + 1, nothing copied) and
                          only computes what follows.

The pages themselves are rows of one pool tensor per layer in HBM (cache.PagedKVCache); a page id indexes the pools of
every layer, and the page tables the attention kernels read are built from the ids handed out here."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from .cache import PagedKVCache

PAGE_TOKENS = 64   # TOKEN_CAPACITY_PER_PAGE (page.hpp:14)


class PageAllocator:
    def __init__(self, num_pages: int, num_heads: int, head_dim: int):
        if num_pages <= 0:
            raise ValueError("PageAllocator must be initialized with num_pages > 0.")     # page_allocator.cpp:21-23
        if num_heads <= 0:
            raise ValueError("num_heads must be positive.")
        if head_dim <= 0:
            raise ValueError("head_dim must be positive.")
        self._L = _lib.load()
        self._h = self._L.pie_page_allocator_create(num_pages, num_heads, head_dim)
        if not self._h:
            raise MemoryError("pie_page_allocator_create failed")
        self.num_heads, self.head_dim = num_heads, head_dim

    def __del__(self):
        if getattr(self, "_h", None):
            self._L.pie_page_allocator_destroy(self._h)
            self._h = None

    def _check(self, page_id: int) -> None:
        if not (0 <= page_id < self.size()):
            raise IndexError(f"Page ID {page_id} is out of range for pool size {self.size()}")   # std::out_of_range, :113-120

    def allocate_page(self):
        """A page id (reference count 1, no tokens), or None when the pool is exhausted (std::nullopt)."""
        i = self._L.pie_page_allocate(self._h)
        return None if i < 0 else int(i)

    def free_page(self, page_id: int) -> None:
        self._check(page_id)
        self._L.pie_page_free(self._h, page_id)

    def add_ref(self, page_id: int) -> None:
        self._check(page_id)
        self._L.pie_page_add_ref(self._h, page_id)

    def get_ref_count(self, page_id: int) -> int:
        self._check(page_id)
        return int(self._L.pie_page_ref_count(self._h, page_id))

    def num_tokens(self, page_id: int) -> int:
        self._check(page_id)
        return int(self._L.pie_page_num_tokens(self._h, page_id))

    def set_num_tokens(self, page_id: int, n: int) -> None:
        self._check(page_id)
        if _lib.check(self._L.pie_page_set_num_tokens(self._h, page_id, n), "set_num_tokens"):
            pass

    def size(self) -> int:
        return int(self._L.pie_page_allocator_size(self._h))

    def get_num_free_pages(self) -> int:
        return int(self._L.pie_page_allocator_num_free(self._h))

    # whole sequences at a time
    def allocate_pages(self, n: int) -> list[int]:
        """n pages, all or nothing; MemoryError when the pool cannot serve them."""
        out = (C.c_int32 * max(n, 1))()
        st = self._L.pie_page_allocate_n(self._h, out, n)
        if st == 7:
            raise MemoryError(f"page pool exhausted: {n} pages wanted, {self.get_num_free_pages()} free")
        _lib.check(st, "allocate_pages")
        return [int(out[i]) for i in range(n)]

    def free_pages(self, ids) -> None:
        arr = (C.c_int32 * max(len(ids), 1))(*ids)
        st = self._L.pie_page_free_n(self._h, arr, len(ids))
        if st == 1:
            raise IndexError("page id out of range")
        _lib.check(st, "free_pages")

    def add_refs(self, ids) -> None:
        arr = (C.c_int32 * max(len(ids), 1))(*ids)
        st = self._L.pie_page_add_ref_n(self._h, arr, len(ids))
        if st == 1:
            raise IndexError("page id out of range")
        _lib.check(st, "add_refs")

    def bench_refcount(self, iterations: int = 20) -> float:
        """ms per iteration of the reference's ReferenceCountingScenario benchmark on this allocator."""
        return float(self._L.pie_page_allocator_bench_refcount(self._h, iterations))

    def stress(self, threads: int, ops_per_thread: int) -> int:
        return int(self._L.pie_page_allocator_stress(self._h, threads, ops_per_thread))


def common_prefix_length(computed_ids, prompt_ids) -> int:
    """PromptCache.reuse_cache's scan (prompt_cache.py:63-67): the shared prefix, never the whole new prompt (its last token
    is always recomputed so that there are logits to sample from)."""
    n = 0
    for i, t in enumerate(computed_ids):
        if i >= len(prompt_ids) - 1 or int(prompt_ids[i]) != int(t):
            break
        n += 1
    return n


class PagedPromptCache:
    """One sequence's KV pages across all layers + the token ids they hold.

        pc = PagedPromptCache(allocator, pools, max_pages)       # pools: [(k_pool, v_pool)] per layer
        rest = pc.reuse_cache(prompt_ids)    # tokens still to run through the model (prompt_cache.py:52-76)
        logits = model(rest[None], cache=pc.cache)
        pc.update(rest)                      # prompt_cache.py:43-50
        other = pc.fork()                    # a second sequence on the same prefix: pages shared by reference count

    Sharing is page-granular: `reuse_cache` keeps the FULL pages of the common prefix (the partially filled page after
    them is released and recomputed -- at most 63 tokens; the reference's contiguous cache keeps the prefix to the token)."""

    def __init__(self, allocator: PageAllocator, pools, max_pages: int, device=None):
        self.alloc, self.pools, self.max_pages = allocator, pools, max_pages
        self.device = device if device is not None else pools[0][0].device
        self.pages: list[int] = []
        self.computed_ids: list[int] = []
        self.table = torch.zeros((1, max_pages), dtype=torch.int32, device=self.device)
        self.cache = [PagedKVCache(k, v, self.table) for (k, v) in pools]

    @property
    def offset(self) -> int:
        return self.cache[0].offset if self.cache else 0

    def _set_offset(self, n: int) -> None:
        for c in self.cache:
            c.offset = n

    def _sync_table(self) -> None:
        row = np.zeros((self.max_pages,), dtype=np.int32)
        row[: len(self.pages)] = self.pages
        self.table.copy_(torch.from_numpy(row)[None])

    def reserve(self, n_tokens: int) -> None:
        """Pages for the first n_tokens tokens of the sequence (ReusableKVCache's capacity growth, reusable.py:60-94)."""
        need = (n_tokens + PAGE_TOKENS - 1) // PAGE_TOKENS
        if need > self.max_pages:
            raise ValueError("sequence longer than the page table")
        if need > len(self.pages):
            self.pages += self.alloc.allocate_pages(need - len(self.pages))
            self._sync_table()

    def update(self, ids) -> None:
        self.computed_ids += [int(t) for t in ids]
        for j, p in enumerate(self.pages):
            self.alloc.set_num_tokens(p, max(0, min(PAGE_TOKENS, len(self.computed_ids) - j * PAGE_TOKENS)))

    def reuse_cache(self, prompt_ids):
        """Returns the part of prompt_ids that still has to go through the model; afterwards the cache holds the shared
        prefix (offset) and has pages for the whole prompt."""
        prompt_ids = [int(t) for t in prompt_ids]
        keep_pages = common_prefix_length(self.computed_ids, prompt_ids) // PAGE_TOKENS
        if len(self.pages) > keep_pages:
            self.alloc.free_pages(self.pages[keep_pages:])
            self.pages = self.pages[:keep_pages]
        kept = keep_pages * PAGE_TOKENS
        self.computed_ids = self.computed_ids[:kept]
        self._set_offset(kept)
        self.reserve(len(prompt_ids))
        self._sync_table()
        return prompt_ids[kept:]

    def fork(self) -> "PagedPromptCache":
        """A second sequence that starts from this one's FULL pages (shared: reference count + 1 each)."""
        other = PagedPromptCache(self.alloc, self.pools, self.max_pages, self.device)
        full = len(self.computed_ids) // PAGE_TOKENS
        self.alloc.add_refs(self.pages[:full])
        other.pages = list(self.pages[:full])
        other.computed_ids = self.computed_ids[: full * PAGE_TOKENS]
        other._set_offset(full * PAGE_TOKENS)
        other._sync_table()
        return other

    def release(self) -> None:
        if self.pages:
            self.alloc.free_pages(self.pages)
        self.pages, self.computed_ids = [], []
        self._set_offset(0)
