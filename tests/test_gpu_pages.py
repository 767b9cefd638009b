"""Prefix reuse over the paged KV cache: a prompt that shares a prefix with a cached one re-uses the
prefix's pages by reference count (PromptCache.reuse_cache, PIE/cache/prompt_cache.py:52-76; ReusableKVCache.reuse,
reusable.py:44-94) and computes only what follows -- same logits as computing the whole prompt on fresh pages."""
import numpy as np
import pytest
import torch

from parity_util import rel_l2, tiny_args

pytestmark = pytest.mark.gpu


def _setup(num_pages=48):
    from proxy_inference_engine_b200.checkpoint import build_model, random_weights
    from proxy_inference_engine_b200.paging import PageAllocator
    args = tiny_args(bits=4)
    model = build_model(args, random_weights(args, seed=3, device="cuda", gen_device="cpu"))
    hkv, d = args.num_key_value_heads, args.head_dim
    pools = [(torch.zeros((num_pages, 64, hkv, d), dtype=torch.bfloat16, device="cuda"),
              torch.zeros((num_pages, 64, hkv, d), dtype=torch.bfloat16, device="cuda")) for _ in range(args.num_hidden_layers)]
    return args, model, pools, PageAllocator(num_pages, hkv, d)


def _run(model, pc, prompt):
    rest = pc.reuse_cache(prompt)
    ids = torch.tensor(rest, dtype=torch.int32, device="cuda")[None]
    logits = model(ids, cache=pc.cache)
    pc.update(rest)
    return rest, logits[0, -1].float().cpu().numpy()


def test_forked_sequence_shares_prefix_pages_and_matches_a_fresh_run():
    from proxy_inference_engine_b200.paging import PagedPromptCache
    args, model, pools, alloc = _setup()
    rng = np.random.default_rng(0)
    prompt_a = rng.integers(0, args.vocab_size, 150).tolist()
    prompt_b = prompt_a[:140] + rng.integers(0, args.vocab_size, 30).tolist()
    a = PagedPromptCache(alloc, pools, max_pages=8)
    rest_a, _ = _run(model, a, prompt_a)
    assert len(rest_a) == 150 and len(a.pages) == 3 and alloc.get_num_free_pages() == 48 - 3
    b = a.fork()                                   # two full pages (128 tokens) shared
    assert b.pages == a.pages[:2] and [alloc.get_ref_count(p) for p in a.pages] == [2, 2, 1]
    rest_b, lb = _run(model, b, prompt_b)
    assert rest_b == prompt_b[128:] and b.offset == 170 and len(b.pages) == 3 and b.pages[:2] == a.pages[:2]
    c = PagedPromptCache(alloc, pools, max_pages=8)
    rest_c, lc = _run(model, c, prompt_b)          # the same prompt on pages of its own
    assert len(rest_c) == 170 and not set(c.pages) & set(a.pages)
    assert rel_l2(lb, lc) < 1e-2 and int(lb.argmax()) == int(lc.argmax())
    # the shared pages hold exactly what a fresh run writes there
    for (kp, vp) in pools:
        for j in range(2):
            assert torch.equal(kp[b.pages[j]], kp[c.pages[j]]) and torch.equal(vp[b.pages[j]], vp[c.pages[j]])
    assert [alloc.num_tokens(p) for p in b.pages] == [64, 64, 42]
    a.release()
    assert [alloc.get_ref_count(p) for p in b.pages[:2]] == [1, 1]      # still owned by b
    b.release(); c.release()
    assert alloc.get_num_free_pages() == 48


def test_reuse_cache_on_the_same_sequence_trims_to_full_pages():
    from proxy_inference_engine_b200.paging import PagedPromptCache
    args, model, pools, alloc = _setup()
    rng = np.random.default_rng(1)
    p1 = rng.integers(0, args.vocab_size, 200).tolist()
    p2 = p1[:100] + rng.integers(0, args.vocab_size, 50).tolist()      # 100 common tokens -> one full page kept
    pc = PagedPromptCache(alloc, pools, max_pages=8)
    _run(model, pc, p1)
    first_page = pc.pages[0]
    rest, l2 = _run(model, pc, p2)
    assert rest == p2[64:] and pc.pages[0] == first_page and len(pc.pages) == 3 and pc.offset == 150
    fresh = PagedPromptCache(alloc, pools, max_pages=8)
    _, lf = _run(model, fresh, p2)
    assert rel_l2(l2, lf) < 1e-2 and int(l2.argmax()) == int(lf.argmax())
    # identical prompt again: everything but the last page's tokens is reused (the last token is always recomputed)
    rest3, l3 = _run(model, pc, p2)
    assert rest3 == p2[128:] and rel_l2(l3, lf) < 1e-2
    pc.release(); fresh.release()
    assert alloc.get_num_free_pages() == 48
    with pytest.raises(ValueError):
        PagedPromptCache(alloc, pools, max_pages=2).reuse_cache(list(range(200)))
    small = __import__("proxy_inference_engine_b200.paging", fromlist=["PageAllocator"]).PageAllocator(2, 2, 64)
    with pytest.raises(MemoryError):
        PagedPromptCache(small, pools, max_pages=8).reuse_cache(list(range(200)))
    assert small.get_num_free_pages() == 2
