"""ctypes binding of libpie_b200.so (include/pie_b200.h).

The library is the product: there is no Python/CPU fallback.  If it is missing, ``load()`` raises and every
operator of this package fails with it.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PIE_LIB") or os.path.join(_HERE, "libpie_b200.so")  # PIE_LIB: experiment builds
CSRC = os.path.join(_HERE, "csrc")

PIE_OK = 0
STATUS = {1: "EINVAL_SHAPE", 2: "EINVAL_DTYPE", 3: "EUNSUPPORTED", 4: "ECUDA", 5: "ENULL", 6: "EALIGN",
          7: "EWORKSPACE", 8: "ESTATIC_LIMIT"}
EPI_STORE, EPI_RESADD, EPI_SILUMUL = 0, 1, 2
DTYPE_BF16, DTYPE_F32 = 0, 1


class QmvSeg(C.Structure):
    _fields_ = [("wq", C.c_void_p), ("scales", C.c_void_p), ("biases", C.c_void_p), ("N", C.c_int)]


class KvLayout(C.Structure):
    _fields_ = [("batch_stride", C.c_longlong), ("head_stride", C.c_longlong), ("tok_stride", C.c_longlong),
                ("page_table", C.c_void_p), ("pt_stride", C.c_int), ("page_size", C.c_int)]


class KvqCache(C.Structure):
    """pie_kvq_cache"""
    _fields_ = [("k_quant", C.c_void_p), ("k_scales", C.c_void_p), ("k_biases", C.c_void_p),
                ("v_quant", C.c_void_p), ("v_scales", C.c_void_p), ("v_biases", C.c_void_p),
                ("capacity", C.c_longlong), ("page_table", C.c_void_p), ("pt_stride", C.c_int), ("page_size", C.c_int),
                ("group_size", C.c_int), ("bits", C.c_int)]


class TpPeers(C.Structure):
    """pie_tp_peers"""
    _fields_ = [("base", C.c_void_p * 8), ("rank", C.c_int), ("world", C.c_int), ("max_elems", C.c_int),
                ("epoch", C.c_void_p)]


class TpFused(C.Structure):
    """pie_tp_fused"""
    _fields_ = [("peers", TpPeers), ("mode", C.c_int), ("h_out", C.c_void_p)]


class StepLayer(C.Structure):
    """pie_step_layer"""
    _fields_ = [("q", QmvSeg), ("k", QmvSeg), ("v", QmvSeg), ("o", QmvSeg), ("gate", QmvSeg), ("up", QmvSeg),
                ("down", QmvSeg), ("input_norm", C.c_void_p), ("post_norm", C.c_void_p), ("kcache", C.c_void_p),
                ("vcache", C.c_void_p), ("kv", KvLayout)]


class StepModel(C.Structure):
    """pie_step_model"""
    _fields_ = [("n_layers", C.c_int), ("hidden", C.c_int), ("ffn", C.c_int), ("Hq", C.c_int), ("Hkv", C.c_int),
                ("D", C.c_int), ("vocab", C.c_int), ("bits", C.c_int), ("group_size", C.c_int),
                ("rms_eps", C.c_float), ("attn_scale", C.c_float), ("rope_freqs", C.c_void_p), ("embed", QmvSeg),
                ("head", QmvSeg), ("final_norm", C.c_void_p)]


class StepIO(C.Structure):
    """pie_step_io"""
    _fields_ = [("B", C.c_int), ("tokens", C.c_void_p), ("positions", C.c_void_p), ("seq", C.c_void_p),
                ("seq_ld", C.c_int), ("h", C.c_void_p), ("qkv", C.c_void_p), ("attn", C.c_void_p),
                ("act", C.c_void_p), ("logits", C.c_void_p), ("attn_ws", C.c_void_p), ("attn_ws_bytes", C.c_size_t),
                ("nsplit", C.c_int), ("qmv_ws", C.c_void_p), ("qmv_ws_bytes", C.c_size_t), ("trace", C.c_void_p)]


_vp, _i, _ll, _f, _sz = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_size_t

# name -> (restype, argtypes); must list every symbol include/pie_b200.h declares (tests check this)
SIGNATURES = {
    "pie_abi_version": (_i, []),
    "pie_status_string": (C.c_char_p, [_i]),
    "pie_last_cuda_error": (C.c_char_p, []),
    "pie_launch_count": (C.c_ulonglong, []),
    "pie_quantize": (_i, [_vp, _i, _vp, _vp, _vp, _ll, _i, _i, _i, _vp]),
    "pie_dequantize": (_i, [_vp, _vp, _vp, _vp, _vp, _ll, _ll, _i, _i, _i, _vp]),
    "pie_qmv_workspace_bytes": (_sz, [_i, _ll, _i, _i, _i, _i]),
    "pie_qmv": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp, _sz, _vp]),
    "pie_qmv_fused": (_i, [_vp, _i, _vp, _i, _i, _i, C.POINTER(QmvSeg), _i, _i, _i, _vp, _f, _i, _vp, _i, _vp,
                           _sz, _i, _vp]),
    "pie_qmm": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp, _sz, _vp]),
    "pie_rms_norm": (_i, [_vp, _vp, _vp, _ll, _i, _f, _vp]),
    "pie_rope": (_i, [_vp, _vp, _ll, _i, _i, _i, _i, _f, _f, _i, _vp, _vp]),
    "pie_rope_kv_append": (_i, [_vp, _vp, _vp, _vp, C.POINTER(KvLayout), _vp, _vp, _i, _i, _i, _i, _vp]),
    "pie_kv_write": (_i, [_vp, _vp, C.POINTER(KvLayout), _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "pie_decode_attn_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "pie_decode_attn_pick_splits": (_i, [_i, _i, _i, _i]),
    "pie_decode_attn": (_i, [_vp, _vp, _vp, _vp, C.POINTER(KvLayout), _vp, _i, _i, _i, _i, _f, _i, _vp, _sz, _vp]),
    "pie_rope_table": (_i, [_vp, _vp, _vp, _i, _i, _vp]),
    "pie_kvq_append": (_i, [_vp, _vp, C.POINTER(KvqCache), _vp, _vp, _i, _i, _i, _i, _vp]),
    "pie_kvq_write": (_i, [C.POINTER(KvqCache), _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "pie_decode_attn_q8_pick_splits": (_i, [_i, _i, _i, _i]),
    "pie_decode_attn_q8": (_i, [_vp, _vp, C.POINTER(KvqCache), _vp, _i, _i, _i, _i, _f, _i, _vp, _sz, _vp]),
    "pie_decode_attn_static": (_i, [_vp, _vp, _vp, _vp, C.POINTER(KvLayout), _vp, _i, _i, _i, _i, _f, _i, _vp]),
    "pie_decode_attn_fused": (_i, [_vp, _vp, _vp, _vp, C.POINTER(KvLayout), _vp, _vp, _vp, _i, _i, _i, _i, _f, _i, _vp,
                                   _sz, _i, _vp]),
    "pie_prefill_attn": (_i, [_vp, _vp, _vp, _vp, C.POINTER(KvLayout), _i, _i, _i, _i, _i, _i, _f, _vp]),
    "pie_silu_mul": (_i, [_vp, _vp, _vp, _ll, _vp]),
    "pie_add": (_i, [_vp, _vp, _vp, _ll, _vp]),
    "pie_argmax": (_i, [_vp, _vp, _i, _i, _vp, _vp, _i, _vp]),
    "pie_qmv_fused_tp": (_i, [_vp, _i, _vp, _i, _i, _i, C.POINTER(QmvSeg), _i, _i, _i, _vp, _f, _i, _vp, _i, _vp,
                              _sz, _i, C.POINTER(TpFused), _vp]),
    "pie_tp_buffer_bytes": (_sz, [_i, _i, _sz]),
    "pie_tp_gather_offset": (_sz, [_i, _i]),
    "pie_tp_allreduce": (_i, [_vp, _vp, _vp, _i, _i, C.POINTER(TpPeers), _vp]),
    "pie_tp_allgather": (_i, [_vp, _i, _i, C.POINTER(TpPeers), _vp]),
    "pie_page_allocator_create": (_vp, [_sz, _i, _i]),
    "pie_page_allocator_destroy": (None, [_vp]),
    "pie_page_allocate": (_ll, [_vp]),
    "pie_page_free": (_i, [_vp, C.c_uint32]),
    "pie_page_add_ref": (_i, [_vp, C.c_uint32]),
    "pie_page_ref_count": (_ll, [_vp, C.c_uint32]),
    "pie_page_num_tokens": (_ll, [_vp, C.c_uint32]),
    "pie_page_set_num_tokens": (_i, [_vp, C.c_uint32, C.c_uint32]),
    "pie_page_allocator_size": (_sz, [_vp]),
    "pie_page_allocator_num_free": (_sz, [_vp]),
    "pie_page_allocate_n": (_i, [_vp, _vp, _i]),
    "pie_page_free_n": (_i, [_vp, _vp, _i]),
    "pie_page_add_ref_n": (_i, [_vp, _vp, _i]),
    "pie_page_allocator_bench_refcount": (C.c_double, [_vp, _i]),
    "pie_page_allocator_stress": (_ll, [_vp, _i, _i]),
    "pie_debug_stamps": (_i, [_vp, _sz]),
    "pie_sample": (_i, [_vp, _vp, _i, _i, _i, _f, _f, _f, _i, _i, _vp, _vp, _vp, _vp, _vp, _i, _vp]),
    "pie_decode_step_plan_bytes": (_sz, [_i]),
    "pie_decode_step_pick_splits": (_i, [_i, _i, _i, _i]),
    "pie_decode_step_attn_ws_bytes": (_sz, [_i, _i, _i, _i]),
    "pie_decode_step_plan": (_i, [C.POINTER(StepModel), C.POINTER(StepLayer), C.POINTER(StepIO), _vp, _sz]),
    "pie_decode_step": (_i, [_vp, _vp, _sz, _vp, _vp]),
}

_lib = None


def build(verbose: bool = False) -> str:
    """Compile csrc/ for sm_100a into libpie_b200.so (nvcc cross-compiles without a GPU)."""
    out = None if verbose else subprocess.DEVNULL
    subprocess.check_call(["make", "-s", "-j8", "-C", CSRC], stdout=out)
    return LIB_PATH


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `make -C {CSRC}` (or __graft_entry__.build()). "
                "proxy_inference_engine_b200 has no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def check(status: int, what: str) -> None:
    """Map a pie_status to the exception the reference raises for the same condition (ops.cpp:82-134)."""
    if status == PIE_OK:
        return
    L = load()
    msg = f"[{what}] {L.pie_status_string(status).decode()} ({STATUS.get(status, status)})"
    if status == 4:
        raise RuntimeError(msg + ": " + L.pie_last_cuda_error().decode())
    raise ValueError(msg)
