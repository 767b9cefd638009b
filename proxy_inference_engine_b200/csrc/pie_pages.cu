// Page allocator of the paged KV cache: fixed pool of 64-token pages, per-page reference counts, lock-free
// LIFO free list.  Host code (the pages themselves live in ONE pool tensor per layer in HBM -- cache.PagedKVCache -- and a
// page id is an index into it; the reference's KVPage owns an MLX array per page, CORE/include/engine/page.hpp:18-123).
//
// Behaviour of CORE/src/engine/page_allocator.cpp:63-158 (allocate_page / free_page / add_ref / get_num_free_pages,
// LIFO order, exhaustion, out-of-range ids), restated over indices instead of node pointers: the free list's head is one
// 64-bit word {tag:16, free count:24, index + 1:24} updated by ONE compare-and-swap per push / pop -- the free count rides
// in the same word (no second atomic per call), and the tag makes a pop immune to the A-B-A interleaving a pointer-only
// Treiber stack is exposed to (pop reads head->next, another thread pops two nodes and pushes the first back, the stale
// next is installed).  Pools of up to 2^24 - 2 pages (a 180 GB GPU holds ~21 000 pages of an 8B model's KV).
#include <atomic>
#include <chrono>
#include <cstdint>
#include <new>
#include <thread>
#include <vector>

#include "../../include/pie_b200.h"

struct pie_page_allocator {
  uint32_t num_pages;
  int num_heads, head_dim;
  std::atomic<uint64_t> head;            // {tag:16, free count:24, index + 1:24}; index field 0 = empty
  std::vector<std::atomic<uint32_t>> ref, ntok;
  std::vector<uint32_t> next;            // index + 1 of the next free page, 0 = end
  explicit pie_page_allocator(uint32_t n) : num_pages(n), head(0), ref(n), ntok(n), next(n) {}
};

namespace {

constexpr uint64_t kIdxMask = 0xffffffull, kCntShift = 24, kCntMask = 0xffffffull, kTagShift = 48;
inline uint64_t pa_word(uint64_t tag, uint64_t count, uint64_t idx1) {
  return (tag & 0xffff) << kTagShift | (count & kCntMask) << kCntShift | (idx1 & kIdxMask);
}
inline void pa_push(pie_page_allocator* a, uint32_t id) {
  uint64_t old = a->head.load(std::memory_order_relaxed);
  for (;;) {
    a->next[id] = (uint32_t)(old & kIdxMask);
    const uint64_t neu = pa_word((old >> kTagShift) + 1, ((old >> kCntShift) & kCntMask) + 1, id + 1);
    if (a->head.compare_exchange_weak(old, neu, std::memory_order_release, std::memory_order_relaxed)) break;
  }
}
inline int64_t pa_pop(pie_page_allocator* a) {
  uint64_t old = a->head.load(std::memory_order_acquire);
  for (;;) {
    const uint32_t top = (uint32_t)(old & kIdxMask);
    if (top == 0) return -1;
    const uint64_t neu = pa_word((old >> kTagShift) + 1, ((old >> kCntShift) & kCntMask) - 1, a->next[top - 1]);
    if (a->head.compare_exchange_weak(old, neu, std::memory_order_acquire, std::memory_order_acquire)) return (int64_t)top - 1;
  }
}

}  // namespace

extern "C" {

pie_page_allocator* pie_page_allocator_create(size_t num_pages, int num_heads, int head_dim) {
  if (num_pages == 0 || num_pages >= 0xffffffull || num_heads <= 0 || head_dim <= 0) return nullptr;  // :20-29 (and the word's 24-bit fields)
  pie_page_allocator* a = new (std::nothrow) pie_page_allocator((uint32_t)num_pages);
  if (!a) return nullptr;
  a->num_heads = num_heads; a->head_dim = head_dim;
  for (uint32_t i = 0; i < a->num_pages; ++i) {   // page 0 on top, as the reference links its nodes (:47-59)
    a->next[i] = i + 1 < a->num_pages ? i + 2 : 0;
    a->ref[i].store(0, std::memory_order_relaxed);
    a->ntok[i].store(0, std::memory_order_relaxed);
  }
  a->head.store(pa_word(0, a->num_pages, 1), std::memory_order_release);
  return a;
}
void pie_page_allocator_destroy(pie_page_allocator* a) { delete a; }

long long pie_page_allocate(pie_page_allocator* a) {
  if (!a) return -1;
  const int64_t id = pa_pop(a);
  if (id < 0) return -1;                                   // pool exhausted (std::nullopt, :65-69)
  a->ref[id].store(1, std::memory_order_release);          // fresh page (:71)
  a->ntok[id].store(0, std::memory_order_release);         // the caller fills tokens (:73)
  return id;
}
int pie_page_free(pie_page_allocator* a, uint32_t id) {
  if (!a) return PIE_ENULL;
  if (id >= a->num_pages) return PIE_EINVAL_SHAPE;         // std::out_of_range (:113-120)
  if (a->ref[id].fetch_sub(1, std::memory_order_acq_rel) == 1) pa_push(a, id);   // :77-83
  return PIE_OK;
}
int pie_page_add_ref(pie_page_allocator* a, uint32_t id) {
  if (!a) return PIE_ENULL;
  if (id >= a->num_pages) return PIE_EINVAL_SHAPE;
  a->ref[id].fetch_add(1, std::memory_order_acq_rel);      // :85-88
  return PIE_OK;
}
long long pie_page_ref_count(const pie_page_allocator* a, uint32_t id) {
  return (!a || id >= a->num_pages) ? -1 : (long long)a->ref[id].load(std::memory_order_acquire);
}
long long pie_page_num_tokens(const pie_page_allocator* a, uint32_t id) {
  return (!a || id >= a->num_pages) ? -1 : (long long)a->ntok[id].load(std::memory_order_acquire);
}
int pie_page_set_num_tokens(pie_page_allocator* a, uint32_t id, uint32_t n) {
  if (!a) return PIE_ENULL;
  if (id >= a->num_pages || n > PIE_PAGE_TOKENS) return PIE_EINVAL_SHAPE;
  a->ntok[id].store(n, std::memory_order_release);
  return PIE_OK;
}
size_t pie_page_allocator_size(const pie_page_allocator* a) { return a ? a->num_pages : 0; }
size_t pie_page_allocator_num_free(const pie_page_allocator* a) {
  return a ? (size_t)((a->head.load(std::memory_order_acquire) >> kCntShift) & kCntMask) : 0;
}

// n pages for one sequence, all or nothing: on exhaustion the pages taken so far go back and PIE_EWORKSPACE is returned
int pie_page_allocate_n(pie_page_allocator* a, int32_t* out, int n) {
  if (!a || (!out && n > 0)) return PIE_ENULL;
  if (n < 0) return PIE_EINVAL_SHAPE;
  for (int i = 0; i < n; ++i) {
    const long long id = pie_page_allocate(a);
    if (id < 0) {
      for (int j = i - 1; j >= 0; --j) pie_page_free(a, (uint32_t)out[j]);
      return PIE_EWORKSPACE;
    }
    out[i] = (int32_t)id;
  }
  return PIE_OK;
}
int pie_page_free_n(pie_page_allocator* a, const int32_t* ids, int n) {
  if (!a || (!ids && n > 0)) return PIE_ENULL;
  int st = PIE_OK;
  for (int i = 0; i < n; ++i) {
    const int s = ids[i] < 0 ? PIE_EINVAL_SHAPE : pie_page_free(a, (uint32_t)ids[i]);
    if (s != PIE_OK) st = s;
  }
  return st;
}
int pie_page_add_ref_n(pie_page_allocator* a, const int32_t* ids, int n) {
  if (!a || (!ids && n > 0)) return PIE_ENULL;
  for (int i = 0; i < n; ++i)
    if (ids[i] < 0 || (uint32_t)ids[i] >= a->num_pages) return PIE_EINVAL_SHAPE;
  for (int i = 0; i < n; ++i) a->ref[ids[i]].fetch_add(1, std::memory_order_acq_rel);
  return PIE_OK;
}

// The reference's BM_PageAllocator_ReferenceCountingScenario (benchmarks/cpp/core/page_allocator_benchmark.cpp:191-300)
// on this allocator: allocate half the pool, add a reference to a quarter, free all once, reallocate what came back,
// release the extra references, final cleanup.  Returns milliseconds per iteration (steady_clock), < 0 on failure.
double pie_page_allocator_bench_refcount(pie_page_allocator* a, int iterations) {
  if (!a || iterations <= 0 || pie_page_allocator_num_free(a) != a->num_pages) return -1.0;
  const size_t n = a->num_pages, p1 = n / 2, p2 = p1 / 2;
  std::vector<uint32_t> pages, again, held;
  pages.reserve(n); again.reserve(n); held.reserve(n);
  const auto t0 = std::chrono::steady_clock::now();
  for (int it = 0; it < iterations; ++it) {
    pages.clear(); again.clear(); held.clear();
    for (size_t i = 0; i < p1; ++i) {
      const long long id = pie_page_allocate(a);
      if (id < 0) return -2.0;
      pages.push_back((uint32_t)id);
    }
    for (size_t i = 0; i < p2; ++i) { pie_page_add_ref(a, pages[i]); again.push_back(pages[i]); }
    for (uint32_t id : pages) pie_page_free(a, id);
    for (size_t i = 0; i < n - p2; ++i) {
      const long long id = pie_page_allocate(a);
      if (id < 0) break;
      held.push_back((uint32_t)id);
    }
    for (uint32_t id : again) pie_page_free(a, id);
    // (the reference's "final cleanup" frees `pages` a second time, underflowing counts of pages it no longer owns; here
    //  the pages taken in the reallocation phase are the ones returned, which leaves the pool whole for the next pass)
    for (uint32_t id : held) pie_page_free(a, id);
  }
  const auto t1 = std::chrono::steady_clock::now();
  if (pie_page_allocator_num_free(a) != a->num_pages) return -3.0;
  return std::chrono::duration<double, std::milli>(t1 - t0).count() / iterations;
}

// tests/cpp/test_page_allocator.cpp:278-330 (HighContentionPushPopStress) with native threads: every thread owns 4 pages
// and keeps swapping one of them for a fresh one.  Returns the number of pages missing from the pool afterwards (0 = no
// leak, no double hand-out), -1 on bad arguments.
long long pie_page_allocator_stress(pie_page_allocator* a, int threads, int ops_per_thread) {
  if (!a || threads <= 0 || ops_per_thread <= 0 || (size_t)threads * 4 > a->num_pages ||
      pie_page_allocator_num_free(a) != a->num_pages)
    return -1;
  std::vector<std::vector<uint32_t>> own(threads);
  for (auto& v : own)
    for (int i = 0; i < 4; ++i) v.push_back((uint32_t)pie_page_allocate(a));
  std::atomic<bool> go{false};
  std::atomic<long long> dup{0};
  std::vector<std::thread> th;
  for (int t = 0; t < threads; ++t) {
    th.emplace_back([&, t] {
      while (!go.load(std::memory_order_acquire)) std::this_thread::yield();
      auto& mine = own[t];
      for (int op = 0; op < ops_per_thread && !mine.empty(); ++op) {
        const size_t i = (size_t)op % mine.size();
        pie_page_free(a, mine[i]);
        long long id = -1;
        for (int r = 0; r < 100 && (id = pie_page_allocate(a)) < 0; ++r) std::this_thread::yield();
        if (id < 0) { mine.erase(mine.begin() + (long)i); continue; }
        if (pie_page_ref_count(a, (uint32_t)id) != 1) dup.fetch_add(1);   // handed out twice
        mine[i] = (uint32_t)id;
      }
    });
  }
  go.store(true, std::memory_order_release);
  for (auto& t : th) t.join();
  for (auto& v : own)
    for (uint32_t id : v) pie_page_free(a, id);
  return (long long)a->num_pages - (long long)pie_page_allocator_num_free(a) + dup.load();
}

}  // extern "C"
