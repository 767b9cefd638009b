"""Ragged step: a mix of prefill chunks and decode rows in ONE model step must give every sequence the
logits it gets when it runs alone through Model.__call__ over its own paged cache (the parity-tested per-sequence path,
PIE/models/llama/language.py:199-210), and the oracle's."""
import numpy as np
import pytest
import torch

from parity_util import make_oracle, rel_l2, tiny_args

pytestmark = pytest.mark.gpu
TOL = 2e-2


def _setup(head_dim=64, n_pages=24):
    from proxy_inference_engine_b200.checkpoint import build_model, random_weights
    from proxy_inference_engine_b200.paging import PageAllocator
    args = tiny_args(bits=4, head_dim=head_dim)
    w = random_weights(args, seed=5, device="cuda", gen_device="cpu")
    model = build_model(args, w)
    hkv, d = args.num_key_value_heads, args.head_dim
    mk = lambda: [(torch.zeros((n_pages, 64, hkv, d), dtype=torch.bfloat16, device="cuda"),
                   torch.zeros((n_pages, 64, hkv, d), dtype=torch.bfloat16, device="cuda"))
                  for _ in range(args.num_hidden_layers)]
    return args, w, model, mk, PageAllocator(n_pages, hkv, d)


def _alone(model, pools, pages, tokens):
    """The per-sequence path: the whole token list through Model.__call__ over a PagedKVCache of its own pages."""
    from proxy_inference_engine_b200.cache import PagedKVCache
    table = torch.tensor([pages], dtype=torch.int32, device="cuda")
    cache = [PagedKVCache(k, v, table) for k, v in pools]
    ids = torch.tensor([tokens], dtype=torch.int32, device="cuda")
    return model(ids, cache=cache)[0].float().cpu().numpy()       # [L, vocab]


@pytest.mark.parametrize("head_dim", [64, 128])
def test_mixed_step_matches_each_sequence_alone_and_the_oracle(head_dim):
    from proxy_inference_engine_b200.ragged import BatchDetails, ragged_step
    args, w, model, mk, alloc = _setup(head_dim)
    rng = np.random.default_rng(0)
    seq = {sid: rng.integers(0, args.vocab_size, n).tolist() for sid, n in ((7, 70), (11, 101), (12, 84))}
    pages = {sid: alloc.allocate_pages((len(t) + 63) // 64) for sid, t in seq.items()}
    pools = mk()
    # step 1: fill the contexts -- 11 prefills 100 tokens, 12 prefills its first 64 (two sequences of unequal length)
    b1 = BatchDetails.build([(11, seq[11][:100], 0, pages[11]), (12, seq[12][:64], 0, pages[12])])
    l1 = ragged_step(model, b1, pools).float().cpu().numpy()
    # step 2: 7 prefills from scratch, 11 decodes one token at position 100, 12 continues with a 20-token chunk
    b2 = BatchDetails.build([(7, seq[7], 0, pages[7]), (11, seq[11][100:], 100, pages[11]),
                             (12, seq[12][64:], 64, pages[12])])
    assert (b2.num_prefill_sequences, b2.num_decode_sequences, b2.total_tokens_in_step) == (2, 1, 91)
    l2 = ragged_step(model, b2, pools, all_logits=True).float().cpu().numpy()
    rows = b2.last_token_rows().tolist()
    ref_pools = mk()
    alone = {sid: _alone(model, ref_pools, pages[sid], seq[sid]) for sid in seq}
    assert rel_l2(l1[0], alone[11][99]) < TOL and rel_l2(l1[1], alone[12][63]) < TOL
    for row, sid in zip(rows, (7, 11, 12)):
        assert rel_l2(l2[row], alone[sid][-1]) < TOL, sid
    # every token of the chunks, not only the last: rows 0..69 are sequence 7, rows 71..90 the chunk of sequence 12
    assert rel_l2(l2[:70], alone[7]) < TOL
    assert rel_l2(l2[71:], alone[12][64:]) < TOL
    # and the CPU oracle on the whole sequences (prefill in one piece)
    o = make_oracle(args, w)
    for row, sid in zip(rows, (7, 11, 12)):
        lo = o(np.asarray([seq[sid]], np.int32), o.make_cache(), last_only=True)[0, -1]
        assert rel_l2(l2[row], lo) < TOL, sid
        assert int(l2[row].argmax()) == int(lo.argmax()) or np.sort(lo)[-1] - np.sort(lo)[-2] < TOL * np.abs(lo).max()
    # the pools hold the same keys as the per-sequence run wrote
    for (k, _), (kr, _) in zip(pools, ref_pools):
        for sid in seq:
            n = len(seq[sid])
            for i, pg in enumerate(pages[sid]):
                m = min(64, n - 64 * i)
                assert rel_l2(k[pg, :m].float().cpu().numpy(), kr[pg, :m].float().cpu().numpy()) < 1e-2


def test_decode_only_batch_with_unequal_contexts():
    """Pure decode step, contexts 5 / 64 / 130: per-sequence positions (the reference's caches assume one offset)."""
    from proxy_inference_engine_b200.ragged import BatchDetails, ragged_step
    args, w, model, mk, alloc = _setup(64)
    rng = np.random.default_rng(1)
    lens = {1: 6, 2: 65, 3: 131}
    seq = {sid: rng.integers(0, args.vocab_size, n).tolist() for sid, n in lens.items()}
    pages = {sid: alloc.allocate_pages((n + 63) // 64) for sid, n in lens.items()}
    pools = mk()
    ragged_step(model, BatchDetails.build([(sid, seq[sid][:-1], 0, pages[sid]) for sid in seq]), pools)
    b = BatchDetails.build([(sid, seq[sid][-1:], lens[sid] - 1, pages[sid]) for sid in seq])
    assert b.num_decode_sequences == 3 and b.total_tokens_in_step == 3
    lg = ragged_step(model, b, pools).float().cpu().numpy()
    ref_pools = mk()
    for i, sid in enumerate(seq):
        assert rel_l2(lg[i], _alone(model, ref_pools, pages[sid], seq[sid])[-1]) < TOL, sid


def test_ragged_step_over_the_quantised_cache():
    """f3 + f4 together: the mixed step over 8-bit pools equals each sequence run alone through Model.__call__ over a
    PagedQuantizedKVCache of its own pages (same kernels for the decode rows, dequantise + prefill kernel for the
    prompts) and the oracle running QuantizedKVCache."""
    from proxy_inference_engine_b200.cache import PagedQuantizedKVCache
    from proxy_inference_engine_b200.ragged import BatchDetails, ragged_step
    args, w, model, mk, alloc = _setup(128)
    hkv, d, n_pages = args.num_key_value_heads, args.head_dim, 24
    dummy = torch.zeros((1, 1), dtype=torch.int32, device="cuda")
    caches = [PagedQuantizedKVCache(n_pages, hkv, d, dummy) for _ in range(args.num_hidden_layers)]
    pools = [(c.keys, c.values) for c in caches]
    rng = np.random.default_rng(2)
    seq = {3: rng.integers(0, args.vocab_size, 70).tolist(), 4: rng.integers(0, args.vocab_size, 40).tolist()}
    pages = {sid: alloc.allocate_pages((len(t) + 63) // 64) for sid, t in seq.items()}
    ragged_step(model, BatchDetails.build([(4, seq[4][:39], 0, pages[4])]), pools)
    b = BatchDetails.build([(3, seq[3], 0, pages[3]), (4, seq[4][39:], 39, pages[4])])
    lg = ragged_step(model, b, pools).float().cpu().numpy()
    o = make_oracle(args, w)
    for i, sid in enumerate((3, 4)):
        lo = o(np.asarray([seq[sid]], np.int32), o.make_cache(kv_bits=8), last_only=True)[0, -1]
        assert rel_l2(lg[i], lo) < TOL, sid
