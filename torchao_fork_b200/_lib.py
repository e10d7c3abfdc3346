"""ctypes binding of libao_b200.so (include/ao_b200.h).

The product path has NO fallback: if the library is missing or a call fails, this raises.
Build with ``python -m torchao_fork_b200.build`` (nvcc, sm_100a)."""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# AO_B200_LIB: another build of the SAME library (A/B measurements, benchmarks/ab_build.sh); never a different implementation
LIB_PATH = os.environ.get("AO_B200_LIB") or os.path.join(_HERE, "lib", "libao_b200.so")

_vp, _i64, _i32, _f32 = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_float



class MegaLayer(ctypes.Structure):
    """ao_b200_mega_layer (include/ao_b200.h)"""
    _fields_ = [(n, _vp) for n in ("wqkv_packed", "wqkv_sz", "wo_packed", "wo_sz", "w13_packed", "w13_sz", "w2_packed",
                                   "w2_sz", "attn_norm", "ffn_norm", "k_cache", "v_cache", "k_pad", "v_pad")]


class MegaModel(ctypes.Structure):
    """ao_b200_mega_model (include/ao_b200.h)"""
    _fields_ = ([(n, ctypes.c_int32) for n in ("n_layer", "dim", "n_head", "n_kv_head", "head_dim", "inter", "vocab",
                                               "group_size", "max_seq", "world", "rank", "n_cta", "timeout_ms",
                                               "attn_blocks_per_cta")]
                + [("eps", _f32), ("reserved", ctypes.c_int32), ("layers", ctypes.POINTER(MegaLayer))]
                + [(n, _vp) for n in ("out_packed", "out_sz", "out_norm", "emb", "freqs", "tok", "pos", "logits", "state")]
                + [("xchg", _vp * 8), ("trace", _vp)])


_mp = ctypes.POINTER(MegaModel)

# name -> argtypes (all return int unless listed in _RESTYPES)
SIGNATURES = {
    "ao_b200_error_string": [_i32],
    "ao_b200_abi_version": [],
    "ao_b200_int4_quantize_pack": [_vp, _i64, _i64, _i64, _i64, _i64, _i32, _f32, _i32, _vp, _vp, _vp],
    "ao_b200_int4_choose_qparams": [_vp, _i64, _i64, _i64, _i32, _f32, _i32, _vp, _vp, _vp],
    "ao_b200_int4_quantize": [_vp, _i64, _i64, _i64, _i32, _vp, _vp, _vp, _vp],
    "ao_b200_int4_pack": [_vp, _i64, _i64, _i64, _i64, _i32, _vp, _vp],
    "ao_b200_int4_unpack": [_vp, _i64, _i64, _vp, _vp],
    "ao_b200_int4_dequant": [_vp, _vp, _i64, _i64, _i32, _vp, _vp],
    "ao_b200_tinygemm_unpack": [_vp, _i64, _i64, _i32, _vp, _vp],
    "ao_b200_tinygemm_dequant": [_vp, _vp, _i64, _i64, _i32, _i32, _vp, _vp],
    "ao_b200_tinygemm_pack": [_vp, _i64, _i64, _i32, _vp, _vp],
    "ao_b200_int4_mm_workspace_bytes": [_i64, _i64, _i64],
    "ao_b200_int4_mm": [_vp, _i64, _i64, _i64, _vp, _vp, _i64, _i64, _i32, _vp, _i64, _i64, _vp, _i64, _vp],
    "ao_b200_int4_gemv_fused": [_vp, _i64, _i64, _i64, _vp, _vp, _i64, _i64, _i32, _vp, _i64, _i64, _vp, _f32, _vp, _i64,
                                _i32, _vp, _i64, _vp],
    "ao_b200_int8_quantize": [_vp, _i64, _i64, _i64, _f32, _i32, _vp, _vp, _vp],
    "ao_b200_int8_pack": [_vp, _i64, _i64, _i64, _i64, _vp, _vp],
    "ao_b200_int8_unpack": [_vp, _i64, _i64, _i64, _i64, _vp, _vp],
    "ao_b200_int8_mm_workspace_bytes": [_i64, _i64, _i64],
    "ao_b200_hqq_blocks": [_i64],
    "ao_b200_hqq_init": [_vp, _i64, _i32, _vp, _vp, _vp],
    "ao_b200_hqq_step": [_vp, _i64, _i32, _vp, _vp, ctypes.c_double, ctypes.c_double, _vp, _vp],
    "ao_b200_hqq_solve": [_vp, _i64, _i32, _vp, _vp, ctypes.c_double, ctypes.c_double, ctypes.c_double, _i32, _vp, _vp, _vp],
    "ao_b200_hqq_finalize": [_vp, _i64, _i32, _vp, _vp, _vp, _vp, _vp, _vp],
    "ao_b200_int8_mm": [_vp, _i64, _i64, _i64, _vp, _vp, _i64, _i64, _vp, _i64, _i64, _vp, _i64, _vp],
    "ao_b200_int8_gemv_fused": [_vp, _i64, _i64, _i64, _vp, _vp, _i64, _i64, _vp, _i64, _i64, _vp, _f32, _vp, _i64, _i32,
                                _vp],
    "ao_b200_rmsnorm": [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _f32, _vp],
    "ao_b200_rope_kv_append": [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _vp],
    "ao_b200_decode_attention_workspace_bytes": [_i32, _i32, _i32],
    "ao_b200_decode_attention": [_vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _f32, _i32, _vp, _i64, _vp],
    "ao_b200_decode_attention_static": [_vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _f32, _i32, _vp],
    "ao_b200_decode_attention_fused_workspace_bytes": [_i32, _i32, _i32, _i32],
    "ao_b200_decode_attention_fused": [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _f32, _i32, _vp, _i64,
                                       _vp],
    "ao_b200_decode_attention_fused_q8": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _f32, _i32,
                                          _vp, _i64, _vp],
    "ao_b200_silu_mul": [_vp, _vp, _i64, _i64, _vp],
    "ao_b200_stream_wait_value32": [_vp, _vp, ctypes.c_uint32],
    "ao_b200_sample": [_vp, _i64, _i64, _i64, _f32, _i32, _vp, _i64, _vp, _vp],
    "ao_b200_tp_buffer_bytes": [_i32, _i64],
    "ao_b200_tp_alloc": [_i64, ctypes.POINTER(ctypes.c_void_p), _vp],
    "ao_b200_tp_open": [_vp, ctypes.POINTER(ctypes.c_void_p)],
    "ao_b200_tp_close": [_vp],
    "ao_b200_tp_free": [_vp],
    "ao_b200_tp_status": [_vp, ctypes.POINTER(ctypes.c_int)],
    "ao_b200_tp_allreduce_add": [_vp, _vp, _i64, ctypes.POINTER(ctypes.c_void_p), _i32, _i32, _i64, _vp],
    "ao_b200_decode_mega_plan_bytes": [_mp],
    "ao_b200_decode_mega_xchg_bytes": [_mp],
    "ao_b200_decode_mega_state_bytes": [],
    "ao_b200_decode_mega_setup": [_mp, _vp, _i64, _vp],
    "ao_b200_decode_mega_step": [_mp, _vp, _vp],
}
_RESTYPES = {"ao_b200_error_string": ctypes.c_char_p, "ao_b200_int4_mm_workspace_bytes": _i64,
             "ao_b200_int8_mm_workspace_bytes": _i64, "ao_b200_hqq_blocks": _i64,
             "ao_b200_decode_attention_workspace_bytes": _i64,
             "ao_b200_decode_attention_fused_workspace_bytes": _i64, "ao_b200_tp_buffer_bytes": _i64,
             "ao_b200_decode_mega_plan_bytes": _i64, "ao_b200_decode_mega_xchg_bytes": _i64,
             "ao_b200_decode_mega_state_bytes": _i64}

DIV_IEEE, DIV_RECIP = 0, 1

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: the B200 CUDA library is not built "
                "(python -m torchao_fork_b200.build). There is no fallback path."
            )
        handle = ctypes.CDLL(LIB_PATH)
        for name, argtypes in SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError if the symbol is not exported
            fn.argtypes = argtypes
            fn.restype = _RESTYPES.get(name, _i32)
        _lib = handle
    return _lib


def check(code, what):
    if code != 0:
        msg = lib().ao_b200_error_string(code).decode()
        raise RuntimeError(f"libao_b200: {what} failed: {msg} (code {code})")
