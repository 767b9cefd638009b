"""CPU: the C-ABI library loads and exports every symbol include/pie_b200.h declares; argument validation and the
pure host-side helpers answer without a GPU (no compute call is made here)."""
import ctypes as C
import os
import re

import pytest

from proxy_inference_engine_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "pie_b200.h")


def _declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(pie_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_path():
    syms = _declared_symbols()
    for must in ("pie_quantize", "pie_dequantize", "pie_qmv", "pie_qmv_fused", "pie_qmm", "pie_rms_norm", "pie_rope",
                 "pie_rope_kv_append", "pie_kv_write", "pie_decode_attn", "pie_prefill_attn", "pie_argmax"):
        assert must in syms


def test_library_exports_every_declared_symbol():
    L = _lib.load()
    for name in _declared_symbols():
        assert hasattr(L, name), f"{name} declared in include/pie_b200.h but not exported by libpie_b200.so"


def test_binding_table_covers_the_header():
    assert sorted(_lib.SIGNATURES) == _declared_symbols()


def test_status_strings_and_version():
    L = _lib.load()
    assert L.pie_abi_version() >= 2
    assert L.pie_status_string(0) == b"ok"
    seen = {L.pie_status_string(i) for i in range(8)}
    assert len(seen) == 8 and b"unknown" not in seen
    assert L.pie_status_string(99) == b"unknown"


def test_validation_happens_before_any_launch():
    """NULL pointers / bad shapes / unsupported quantisation come back as status codes (the Python layer turns them
    into ValueError like MLX/mlx/ops.cpp:82-134); none of these reaches a kernel launch."""
    L = _lib.load()
    n0 = L.pie_launch_count()
    fake = C.c_void_p(0x1000)  # never dereferenced: validation fails first
    assert L.pie_qmv(None, None, None, None, None, 1, 64, 64, 64, 4, None, 0, None) == 5          # ENULL
    assert L.pie_qmv(fake, fake, fake, fake, fake, 1, 64, 64, 48, 4, None, 0, None) == 3          # EUNSUPPORTED gs
    assert L.pie_qmv(fake, fake, fake, fake, fake, 1, 64, 64, 64, 5, None, 0, None) == 3          # EUNSUPPORTED bits
    assert L.pie_qmv(fake, fake, fake, fake, fake, -1, 64, 64, 64, 4, None, 0, None) == 1         # EINVAL_SHAPE
    assert L.pie_qmv(fake, fake, fake, fake, fake, 0, 64, 64, 64, 4, None, 0, None) == 0          # M == 0 is a no-op
    assert L.pie_quantize(fake, 7, fake, fake, fake, 4, 64, 64, 4, None) == 2                     # EINVAL_DTYPE
    assert L.pie_quantize(fake, 0, fake, fake, fake, 4, 100, 64, 4, None) == 1                    # K % gs
    assert L.pie_rms_norm(fake, fake, fake, 0, 64, 1e-5, None) == 0                               # empty input
    assert L.pie_rope(fake, fake, 1, 1, 64, 65, 0, 1e4, 1.0, 0, None, None) == 1                  # dims > D
    lay = _lib.KvLayout(0, 64, 60, None, 0, 0)                                                    # tok_stride < D
    assert L.pie_decode_attn(fake, fake, fake, fake, C.byref(lay), fake, 1, 4, 2, 64, 0.125, 1, None, 0, None) == 1
    lay = _lib.KvLayout(0, 64, 64, None, 0, 0)
    assert L.pie_decode_attn(fake, fake, fake, fake, C.byref(lay), fake, 1, 4, 3, 64, 0.125, 1, None, 0, None) == 1
    assert L.pie_decode_attn(fake, fake, fake, fake, C.byref(lay), fake, 1, 4, 2, 96, 0.125, 1, None, 0, None) == 3
    assert L.pie_decode_attn(fake, fake, fake, fake, C.byref(lay), fake, 1, 4, 2, 64, 0.125, 4, None, 0, None) == 7
    assert L.pie_launch_count() == n0


def test_check_maps_status_to_reference_exceptions():
    with pytest.raises(ValueError):
        _lib.check(1, "quantized_matmul")
    with pytest.raises(ValueError):
        _lib.check(3, "quantized_matmul")
    with pytest.raises(RuntimeError):
        _lib.check(4, "quantized_matmul")
    _lib.check(0, "ok")


def test_workspace_sizing_is_pure_host_logic():
    L = _lib.load()
    # 8B shapes: workspace is only needed when the plan splits K
    for (n, k, dual) in ((6144, 4096, 0), (4096, 4096, 0), (14336, 4096, 1), (4096, 14336, 0), (128256, 4096, 0)):
        nb = L.pie_qmv_workspace_bytes(1, n, k, 4, dual, 0)
        assert nb % 4 == 0 and nb < (64 << 20)
    assert L.pie_qmv_workspace_bytes(1, 4096, 4096, 3, 0, 0) == 0      # unsupported bits -> 0
    assert L.pie_qmv_workspace_bytes(1, 4096, 4100, 4, 0, 0) == 0      # K off the fast path -> 0
    assert L.pie_decode_attn_workspace_bytes(1, 32, 128, 1) > 0
    assert L.pie_decode_attn_workspace_bytes(0, 32, 128, 4) == 0
    a = L.pie_decode_attn_workspace_bytes(2, 32, 128, 8)
    assert a >= 2 * 32 * 8 * (128 + 2) * 4


def test_pick_splits_fills_the_machine_and_respects_context():
    L = _lib.load()
    assert L.pie_decode_attn_pick_splits(1, 32, 8, 1) == 1
    assert L.pie_decode_attn_pick_splits(1, 32, 8, 64) == 1
    s = L.pie_decode_attn_pick_splits(1, 32, 8, 8192)
    assert 1 < s <= 64
    assert L.pie_decode_attn_pick_splits(32, 32, 8, 2048) <= s
    assert L.pie_decode_attn_pick_splits(0, 32, 8, 2048) == 1
    # one wave of resident CTAs (2 per SM on 148 SMs), never a wave and a fraction: measured optimum on B200
    assert L.pie_decode_attn_pick_splits(32, 32, 8, 2048) == 1      # 256 CTAs per split
    assert L.pie_decode_attn_pick_splits(16, 32, 8, 4096) == 2
    assert L.pie_decode_attn_pick_splits(8, 32, 8, 2048) == 4
    assert L.pie_decode_attn_pick_splits(64, 32, 8, 2048) == 1      # more than a wave already
    # quantised-KV kernels: same rule, and at least 512 tokens per split
    assert L.pie_decode_attn_q8_pick_splits(32, 32, 8, 2048) == 1
    assert L.pie_decode_attn_q8_pick_splits(16, 32, 8, 4096) == 2
    assert L.pie_decode_attn_q8_pick_splits(1, 32, 8, 8192) == 16
    assert L.pie_decode_attn_q8_pick_splits(1, 32, 8, 100) == 1


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    """No CPU fallback: without libpie_b200.so the product path raises."""
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "libpie_b200.so"))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _lib.load()
