"""pie_decode_attn_static (include/pie_b200.h): the split-K flash-decode with its scratch in a static device array, for
callers that cannot allocate.  Called through the C ABI, compared with pie_decode_attn (caller-owned workspace) on two
calls in a row; the file sorts last on purpose."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

STATIC_BYTES = 4 << 20   # PIE_STATIC_SCRATCH_BYTES


def _case(seed, B, Hq, Hkv, D, T, pos):
    g = torch.Generator(device="cuda").manual_seed(seed)
    q = torch.randn((B, Hq, D), generator=g, device="cuda").to(torch.bfloat16)
    k = torch.randn((B, Hkv, T, D), generator=g, device="cuda").to(torch.bfloat16)
    v = torch.randn((B, Hkv, T, D), generator=g, device="cuda").to(torch.bfloat16)
    positions = torch.tensor(pos, dtype=torch.int32, device="cuda")
    return q, k, v, positions


def test_static_scratch_entry_matches_the_workspace_entry_on_two_calls_in_a_row():
    from proxy_inference_engine_b200 import _lib
    from proxy_inference_engine_b200._lib import KvLayout
    L = _lib.load()
    st = torch.cuda.current_stream().cuda_stream
    for call, (B, Hq, Hkv, D, T, nsplit, pos) in enumerate([(2, 32, 8, 128, 700, 8, [650, 37]),
                                                           (3, 8, 2, 64, 300, 4, [299, 0, 128])]):
        q, k, v, positions = _case(call, B, Hq, Hkv, D, T, pos)
        kv = KvLayout(Hkv * T * D, T * D, D, None, 0, 0)
        nb = L.pie_decode_attn_workspace_bytes(B, Hq, D, nsplit)
        assert 0 < nb <= STATIC_BYTES
        ws = torch.zeros(nb, dtype=torch.uint8, device="cuda")
        ref = torch.empty((B, Hq * D), dtype=torch.bfloat16, device="cuda")
        got = torch.full((B, Hq * D), float("nan"), dtype=torch.bfloat16, device="cuda")
        scale = D ** -0.5
        _lib.check(L.pie_decode_attn(ref.data_ptr(), q.data_ptr(), k.data_ptr(), v.data_ptr(), C.byref(kv),
                                     positions.data_ptr(), B, Hq, Hkv, D, scale, nsplit, ws.data_ptr(), nb, st), "decode_attn")
        before = torch.cuda.memory_allocated()
        for _ in range(2):   # the second call finds the first one's partials in the array
            _lib.check(L.pie_decode_attn_static(got.data_ptr(), q.data_ptr(), k.data_ptr(), v.data_ptr(), C.byref(kv),
                                                positions.data_ptr(), B, Hq, Hkv, D, scale, nsplit, st), "decode_attn_static")
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        r, g_ = ref.float(), got.float()
        assert torch.isfinite(g_).all()
        # same kernel, same split order: identical in practice; the bound is the bf16 tolerance of the attention tests
        assert (g_ - r).abs().max().item() <= 2e-2 * max(1.0, r.abs().max().item())
        assert torch.equal(got, ref)


def test_static_scratch_entry_rejects_requests_beyond_its_array():
    from proxy_inference_engine_b200 import _lib
    from proxy_inference_engine_b200._lib import KvLayout
    L = _lib.load()
    B, Hq, Hkv, D, T, nsplit = 64, 32, 8, 128, 64, 16      # 64 * 32 * 16 * 130 * 4 B = 17 MB of partials
    assert L.pie_decode_attn_workspace_bytes(B, Hq, D, nsplit) > STATIC_BYTES
    q, k, v, positions = _case(9, 1, Hq, Hkv, D, T, [3])   # (never read: the call is refused before any launch)
    kv = KvLayout(Hkv * T * D, T * D, D, None, 0, 0)
    out = torch.zeros((1, Hq * D), dtype=torch.bfloat16, device="cuda")
    n0 = L.pie_launch_count()
    status = L.pie_decode_attn_static(out.data_ptr(), q.data_ptr(), k.data_ptr(), v.data_ptr(), C.byref(kv),
                                      positions.data_ptr(), B, Hq, Hkv, D, D ** -0.5, nsplit,
                                      torch.cuda.current_stream().cuda_stream)
    assert status == 8                                        # PIE_ESTATIC_LIMIT
    assert b"static scratch" in L.pie_status_string(status)
    assert L.pie_launch_count() == n0
    with pytest.raises(ValueError):
        _lib.check(status, "decode_attn_static")
