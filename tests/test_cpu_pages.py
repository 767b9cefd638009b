"""Page allocator -- the reference's own test cases (tests/cpp/test_page_allocator.cpp:60-330) on
csrc/pie_pages.cu through the C ABI, plus prefix-length logic of PromptCache.reuse_cache (prompt_cache.py:52-76).
Host code: runs without a GPU."""
import threading

import pytest

from proxy_inference_engine_b200.paging import PageAllocator, common_prefix_length

HEADS, DIM = 8, 128


def test_constructor_arguments():
    PageAllocator(4, HEADS, DIM)                                   # ConstructorValidArgs
    for bad in ((0, HEADS, DIM), (4, 0, DIM), (4, HEADS, 0)):      # ConstructorInvalidArgs
        with pytest.raises(ValueError):
            PageAllocator(*bad)


def test_exhaust_and_refill():
    a = PageAllocator(4, HEADS, DIM)
    assert a.size() == 4 and a.get_num_free_pages() == 4
    ids = [a.allocate_page() for _ in range(4)]
    assert sorted(ids) == [0, 1, 2, 3] and a.get_num_free_pages() == 0
    assert a.allocate_page() is None                               # ExhaustionReturnsNullopt
    for i in ids:
        a.free_page(i)
    assert a.get_num_free_pages() == 4
    assert a.allocate_page() is not None and a.get_num_free_pages() == 3


def test_single_page_pool():
    a = PageAllocator(1, HEADS, DIM)
    i = a.allocate_page()
    assert i == 0 and a.allocate_page() is None
    a.free_page(i)
    assert a.get_num_free_pages() == 1 and a.allocate_page() == 0


def test_lifo_order():
    a = PageAllocator(16, HEADS, DIM)
    first = [a.allocate_page() for _ in range(16)]
    for i in reversed(first):
        a.free_page(i)
    assert [a.allocate_page() for _ in range(16)] == first and a.get_num_free_pages() == 0


def test_reference_counting_and_tokens():
    a = PageAllocator(1, HEADS, DIM)
    i = a.allocate_page()
    assert a.get_ref_count(i) == 1 and a.num_tokens(i) == 0
    a.add_ref(i); a.add_ref(i)
    assert a.get_ref_count(i) == 3
    a.free_page(i); a.free_page(i)
    assert a.get_ref_count(i) == 1 and a.get_num_free_pages() == 0
    a.set_num_tokens(i, 64)
    assert a.num_tokens(i) == 64
    a.free_page(i)
    assert a.get_num_free_pages() == 1
    assert a.num_tokens(a.allocate_page()) == 0                    # a fresh page starts empty


def test_invalid_ids_raise():
    a = PageAllocator(5, HEADS, DIM)
    for call in (a.get_ref_count, a.free_page, a.add_ref):
        for bad in (5, 100, 999):
            with pytest.raises(IndexError):
                call(bad)
    with pytest.raises(IndexError):
        a.free_pages([0, 7])


def test_bulk_allocation_is_all_or_nothing():
    a = PageAllocator(8, HEADS, DIM)
    got = a.allocate_pages(5)
    assert len(set(got)) == 5 and a.get_num_free_pages() == 3
    with pytest.raises(MemoryError):
        a.allocate_pages(4)
    assert a.get_num_free_pages() == 3                              # nothing kept from the failed request
    a.add_refs(got[:2])
    a.free_pages(got)
    assert a.get_num_free_pages() == 6 and a.get_ref_count(got[0]) == 1
    a.free_pages(got[:2])
    assert a.get_num_free_pages() == 8


def test_concurrent_producers_and_consumer():
    """ConcurrentAllocFreeProducersConsumer: producers free pages while a consumer allocates all of them (ctypes releases
    the GIL during the calls, so the threads do interleave inside the allocator)."""
    producers, per = 7, 512
    a = PageAllocator(producers * per, HEADS, DIM)
    initial = [a.allocate_pages(per) for _ in range(producers)]
    assert a.get_num_free_pages() == 0
    got, start = [], threading.Event()

    def consumer():
        start.wait()
        while len(got) < producers * per:
            i = a.allocate_page()
            if i is not None:
                got.append(i)

    def producer(p):
        start.wait()
        for i in initial[p]:
            a.free_page(i)

    th = [threading.Thread(target=consumer)] + [threading.Thread(target=producer, args=(p,)) for p in range(producers)]
    for t in th:
        t.start()
    start.set()
    for t in th:
        t.join(timeout=60)
    assert len(got) == producers * per == len(set(got)) and a.get_num_free_pages() == 0


def test_concurrent_free_of_a_shared_page():
    a = PageAllocator(1, HEADS, DIM)
    i = a.allocate_page()
    for _ in range(9):
        a.add_ref(i)
    th = [threading.Thread(target=a.free_page, args=(i,)) for _ in range(10)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert a.get_num_free_pages() == 1 and a.allocate_page() == i and a.get_ref_count(i) == 1


def test_high_contention_stress_native_threads():
    """HighContentionPushPopStress with C++ threads inside the library: no page lost, none handed out twice."""
    a = PageAllocator(128, HEADS, DIM)
    assert a.stress(threads=16, ops_per_thread=20000) == 0
    assert a.get_num_free_pages() == 128


def test_reference_scenario_benchmark_runs():
    a = PageAllocator(10000, 40, 128)
    ms = a.bench_refcount(5)
    assert 0 < ms < 50 and a.get_num_free_pages() == 10000


def test_common_prefix_length_follows_reuse_cache():
    assert common_prefix_length([], [1, 2, 3]) == 0
    assert common_prefix_length([1, 2, 3, 4], [1, 2, 9, 4]) == 2
    assert common_prefix_length([1, 2, 3], [1, 2, 3]) == 2             # the last prompt token is always recomputed
    assert common_prefix_length([1, 2, 3, 4, 5], [1, 2, 3]) == 2
    assert common_prefix_length([1, 2], [1, 2, 3, 4]) == 2
