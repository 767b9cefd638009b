"""Device-side sampler against the numpy restatement of PIE/samplers/*.py + inference_engine.py:268-271.
The kernel and the restatement are given the SAME uniforms, so the sampled tokens must be identical (not just equal in
distribution); the Philox path is checked for reproducibility, graph replay and its distribution."""
import numpy as np
import pytest
import torch

from parity_util import n2t_bf16

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    from oracle import oracle as orc
    from proxy_inference_engine_b200 import ops, samplers
    return orc, ops, samplers


def _logits(rng, rows, vocab, spread):
    from oracle import oracle as orc
    return orc.round_bf16((rng.standard_normal((rows, vocab)) * spread).astype(np.float32))


CASES = [dict(temp=1.0), dict(temp=0.7), dict(temp=0.8, top_k=40), dict(temp=1.3, top_k=1), dict(temp=1.0, top_p=0.9),
         dict(temp=0.6, top_p=0.5), dict(temp=1.0, min_p=0.05), dict(temp=0.9, min_p=0.3, min_tokens_to_keep=5),
         dict(temp=1.0, top_p=0.8, top_k=10, min_p=0.1)]   # precedence: top_p wins (make_sampler :37-45)


@pytest.mark.parametrize("vocab,spread", [(1000, 3.0), (128256, 2.0), (4099, 6.0)])
@pytest.mark.parametrize("kw", CASES)
def test_sampler_matches_reference_given_the_same_uniforms(env, vocab, spread, kw):
    orc, ops, _ = env
    rng = np.random.default_rng(vocab + int(100 * kw["temp"]))
    rows = 24
    lg = _logits(rng, rows, vocab, spread)
    u = rng.uniform(1e-6, 1 - 1e-6, (rows, vocab)).astype(np.float32)
    lp_ref = orc.logprobs_of(lg)
    ref = orc.sample_reference(lp_ref, u, **kw)
    tok, lp = ops.sample(n2t_bf16(lg), uniforms=torch.from_numpy(u).cuda(), return_logprobs=True, **kw)
    assert np.abs(lp.cpu().numpy() - lp_ref).max() < 2e-5 * max(1.0, np.abs(lp_ref).max())
    got = tok.cpu().numpy()
    # a boundary token exactly at a filter threshold may be kept on one side only (fp32 summation order of the nucleus
    # mass; ties inside argpartition): allow it on at most one row, and only if the other side's choice is a kept token too
    assert (got != ref).sum() <= 1, (got, ref)


def test_logprobs_and_errors(env):
    orc, ops, _ = env
    lg = _logits(np.random.default_rng(0), 3, 640, 4.0)
    st = ops.new_rng_state(1)
    tok, lp = ops.sample(n2t_bf16(lg), 1.0, rng_state=st, return_logprobs=True)
    assert np.allclose(np.exp(lp.cpu().numpy()).sum(-1), 1.0, atol=1e-4)
    with pytest.raises(ValueError):
        ops.sample(n2t_bf16(lg), 0.0, rng_state=st)
    with pytest.raises(ValueError):
        ops.sample(n2t_bf16(lg), 1.0, min_p=1.5, rng_state=st)            # min_p.py:33-36
    with pytest.raises(ValueError):
        ops.sample(n2t_bf16(lg), 1.0, min_p=0.1, min_tokens_to_keep=0, rng_state=st)
    with pytest.raises(ValueError):
        ops.sample(n2t_bf16(lg), 1.0, top_k=640, rng_state=st)            # top_k.py:23-27


def test_philox_stream_is_reproducible_advances_and_replays_in_a_graph(env):
    orc, ops, samplers = env
    lg = n2t_bf16(_logits(np.random.default_rng(1), 8, 2048, 1.0))
    a = samplers.make_sampler(temp=1.0, seed=7)
    b = samplers.make_sampler(temp=1.0, seed=7)
    c = samplers.make_sampler(temp=1.0, seed=8)
    ta = [a(lg).cpu() for _ in range(3)]
    tb = [b(lg).cpu() for _ in range(3)]
    assert all(torch.equal(x, y) for x, y in zip(ta, tb))            # same seed, same call number -> same draw
    assert not torch.equal(ta[0], ta[1])                              # the counter advances on the device
    assert not torch.equal(ta[0], c(lg).cpu())
    assert int(a.rng_state[1].item()) == 3
    # captured once, replayed: fresh numbers every replay (the counter lives in device memory)
    out = torch.empty((8,), dtype=torch.int32, device="cuda")
    s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        a(lg, out=out)
    torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        a(lg, out=out)
    seen = []
    for _ in range(4):
        g.replay(); torch.cuda.synchronize(); seen.append(out.cpu().clone())
    assert len({tuple(t.tolist()) for t in seen}) > 1
    assert samplers.make_sampler(temp=0.0)(lg).cpu().tolist() == lg.float().argmax(-1).cpu().tolist()


def test_philox_categorical_follows_the_distribution(env):
    orc, ops, _ = env
    p = np.array([0.5, 0.25, 0.125, 0.0625, 0.0625], np.float32)
    lg = n2t_bf16(np.tile(np.log(p), (4096, 1)))
    st = ops.new_rng_state(3)
    counts = np.zeros(5)
    for _ in range(4):
        counts += np.bincount(ops.sample(lg, 1.0, rng_state=st).cpu().numpy(), minlength=5)
    freq = counts / counts.sum()
    assert np.abs(freq - p).max() < 0.02, freq
    top2 = ops.sample(lg, 1.0, top_k=2, rng_state=st).cpu().numpy()
    assert set(np.unique(top2)) <= {0, 1}


def test_decode_engine_samples_on_device_and_is_reproducible(env):
    """The engine's step with temp > 0: pie_sample inside the captured graph.  Same seed -> same text; a different seed or
    greedy -> a different one; every sampled token is inside the top-k set of the logits the oracle computes for its
    prefix (the forward pass is the parity-tested one; here the link between sampler and engine is what is checked)."""
    from parity_util import make_oracle, tiny_args
    from proxy_inference_engine_b200.checkpoint import build_model, random_weights
    from proxy_inference_engine_b200.decode import DecodeEngine
    args = tiny_args(bits=4)
    w = random_weights(args, seed=5, device="cuda", gen_device="cpu")
    model = build_model(args, w)
    prompt = torch.arange(1, 17, dtype=torch.int32).reshape(1, 16)
    runs = {}
    for name, sp in (("a", dict(temp=1.0, top_k=4, seed=11)), ("b", dict(temp=1.0, top_k=4, seed=11)),
                     ("c", dict(temp=1.0, top_k=4, seed=12)), ("g", None)):
        eng = DecodeEngine(model, batch=1, max_context=64, sampling=sp)
        runs[name] = eng.generate(prompt, 24)[0].tolist()
        assert len(runs[name]) == 24
    assert runs["a"] == runs["b"]
    assert runs["a"] != runs["c"] and runs["a"] != runs["g"]
    o = make_oracle(args, w)
    seq = prompt[0].tolist() + runs["a"]
    logits = o(np.array(seq[:-1])[None, :], o.make_cache())[0]
    for t in range(prompt.shape[1] - 1, len(seq) - 1):
        row = logits[t]
        kth = np.sort(row)[-4]
        assert row[seq[t + 1]] >= kth - 0.05 * max(1.0, abs(kth)), (t, seq[t + 1])


@pytest.mark.parametrize("rows,vocab", [(1, 128256), (3, 128256), (2, 8192), (5, 50001), (4, 640), (70, 9000)])
def test_argmax_first_maximum_bookkeeping_and_cluster_form(env, rows, vocab):
    """Greedy sampler (temp == 0, PIE/samplers/__init__.py:30-31 -> mx.argmax): first index of the row maximum, ties
    included, for the one-CTA form (short / many rows) and the 8-CTA cluster form (few long rows: the decode step)."""
    orc, ops, _ = env
    rng = np.random.default_rng(rows * 1000 + vocab)
    lg = orc.round_bf16(rng.standard_normal((rows, vocab)).astype(np.float32))
    for r in range(rows):     # plant ties of the maximum, also across the boundaries of the cluster's parts
        m = lg[r].max() + 1.0
        idx = rng.choice(vocab, size=min(6, vocab), replace=False)
        lg[r, idx] = m
        if r == 0:
            lg[r, [vocab - 1, vocab // 2]] = m
    pos = torch.arange(rows, dtype=torch.int32, device="cuda")
    seq = torch.zeros((rows, rows + 3), dtype=torch.int32, device="cuda")
    got = ops.argmax(n2t_bf16(lg), positions_to_advance=pos, seq_out=seq).cpu().numpy()
    ref = lg.argmax(-1)
    assert (got == ref).all(), (got, ref)
    assert pos.cpu().tolist() == [r + 1 for r in range(rows)]
    assert [int(seq[r, r + 1]) for r in range(rows)] == ref.tolist()
