"""ctypes binding of libfa_b200.so (C ABI in include/fa_b200.h).

The library is torch-free; this module only mirrors ``FaFwdParams`` and loads the shared object
that build.sh / __graft_entry__.build() put next to this file.  There is no fallback: if the
library is missing, importing the package fails loudly (north_star: "no CPU fallback").
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FA_B200_LIB", os.path.join(_HERE, "libfa_b200.so"))

FA_OK = 0
FA_ERR_WORKSPACE = -4
ABI_VERSION = 5
FA_KERNEL_NONE, FA_KERNEL_TC5_PREFILL, FA_KERNEL_TILE, FA_KERNEL_DECODE = 0, 1, 2, 3
OP_FWD, OP_VARLEN, OP_KVCACHE = 0, 1, 2

_vp, _i64, _i32, _f32 = C.c_void_p, C.c_int64, C.c_int32, C.c_float


class FaFwdParams(C.Structure):
    """Field-for-field mirror of ``struct FaFwdParams`` (include/fa_b200.h)."""
    _fields_ = [
        ("q", _vp), ("k", _vp), ("v", _vp), ("o", _vp), ("softmax_lse", _vp),
        ("q_batch_stride", _i64), ("q_row_stride", _i64), ("q_head_stride", _i64),
        ("k_batch_stride", _i64), ("k_row_stride", _i64), ("k_head_stride", _i64),
        ("v_batch_stride", _i64), ("v_row_stride", _i64), ("v_head_stride", _i64),
        ("o_batch_stride", _i64), ("o_row_stride", _i64), ("o_head_stride", _i64),
        ("b", _i32), ("seqlen_q", _i32), ("seqlen_k", _i32), ("h", _i32), ("h_k", _i32), ("d", _i32),
        ("is_bf16", _i32), ("softmax_scale", _f32),
        ("is_causal", _i32), ("window_left", _i32), ("window_right", _i32),
        ("alibi_slopes", _vp), ("alibi_slopes_batch_stride", _i64),
        ("cu_seqlens_q", _vp), ("cu_seqlens_k", _vp), ("seqused_k", _vp),
        ("is_seqlens_k_cumulative", _i32),
        ("block_table", _vp), ("block_table_batch_stride", _i64),
        ("page_block_size", _i32), ("max_blocks_per_seq", _i32), ("num_blocks", _i32),
        ("cache_batch_idx", _vp), ("total_q", _i64), ("total_k", _i64),
        ("knew", _vp), ("vnew", _vp),
        ("knew_batch_stride", _i64), ("knew_row_stride", _i64), ("knew_head_stride", _i64),
        ("vnew_batch_stride", _i64), ("vnew_row_stride", _i64), ("vnew_head_stride", _i64),
        ("seqlen_knew", _i32),
        ("rotary_cos", _vp), ("rotary_sin", _vp),
        ("rotary_dim", _i32), ("seqlen_ro", _i32), ("is_rotary_interleaved", _i32),
        ("num_splits", _i32), ("workspace", _vp), ("workspace_bytes", C.c_size_t), ("partial_out", _i32), ("kv_shard_start", _i32), ("kv_shard_len", _i32),
        ("num_sms", _i32), ("force_kernel", _i32), ("kernel_used", _i32), ("launches", _i32),
        ("workspace_clean", _i32), ("fused_exchange", _vp),
    ]


EXPORTS = (
    "fa_b200_abi_version", "fa_b200_params_size", "fa_b200_last_error",
    "fa_b200_num_splits_heuristic", "fa_b200_workspace_bytes",
    "fa_b200_fwd", "fa_b200_varlen_fwd", "fa_b200_fwd_kvcache", "fa_b200_combine",
    "fa_b200_append_kv", "fa_b200_exchange_combine", "fa_b200_exchange_params_size",
    "fa_b200_fwd_kvcache_static_scratch", "fa_b200_fused_exchange_size", "fa_b200_fused_append_supported",
    "fa_b200_reshape_and_cache",
)
STATIC_SCRATCH_BYTES = 4 << 20      # FA_B200_STATIC_SCRATCH_BYTES (include/fa_b200.h)


FA_MAX_PEERS = 8


class FaExchangeParams(C.Structure):
    """Mirror of FaExchangeParams (include/fa_b200.h)."""
    _fields_ = [("recv", _vp * FA_MAX_PEERS), ("flags", _vp * FA_MAX_PEERS), ("local_record", _vp), ("state", _vp),
                ("rank", _i32), ("world", _i32), ("b", _i32), ("h", _i32), ("seqlen_q", _i32), ("d", _i32),
                ("record_elems", _i64), ("o", _vp), ("softmax_lse", _vp),
                ("o_batch_stride", _i64), ("o_row_stride", _i64), ("o_head_stride", _i64), ("out_dtype", _i32)]


class FaFusedExchange(C.Structure):
    """Mirror of FaFusedExchange (include/fa_b200.h): sequence-sharded decode in one kernel per rank."""
    _fields_ = [("recv", _vp * FA_MAX_PEERS), ("flags", _vp * FA_MAX_PEERS), ("state", _vp), ("split_counters", _vp),
                ("rank", _i32), ("world", _i32), ("max_units", _i32), ("unit_elems", _i64)]


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with vllm-flash-attention_b200/build.sh "
            "(or __graft_entry__.build()). There is no fallback path.")
    lib = C.CDLL(LIB_PATH)
    lib.fa_b200_abi_version.restype = C.c_int
    lib.fa_b200_params_size.restype = C.c_size_t
    lib.fa_b200_last_error.restype = C.c_char_p
    lib.fa_b200_num_splits_heuristic.restype = C.c_int
    lib.fa_b200_num_splits_heuristic.argtypes = [C.c_int] * 4
    lib.fa_b200_workspace_bytes.restype = C.c_size_t
    lib.fa_b200_workspace_bytes.argtypes = [C.POINTER(FaFwdParams), C.c_int]
    for name in ("fa_b200_fwd", "fa_b200_varlen_fwd", "fa_b200_fwd_kvcache", "fa_b200_append_kv",
                 "fa_b200_fwd_kvcache_static_scratch"):
        fn = getattr(lib, name)
        fn.restype = C.c_int
        fn.argtypes = [C.POINTER(FaFwdParams), _vp]
    lib.fa_b200_fused_append_supported.restype = C.c_int
    lib.fa_b200_fused_append_supported.argtypes = [C.POINTER(FaFwdParams)]
    lib.fa_b200_reshape_and_cache.restype = C.c_int
    lib.fa_b200_reshape_and_cache.argtypes = [_vp] * 5 + [_i32] * 4 + [_i64] * 11 + [_vp]
    lib.fa_b200_combine.restype = C.c_int
    lib.fa_b200_combine.argtypes = [_vp, _vp, _vp, _vp, _i32, _i64, _i32, _i64, _i64, _i32, _i32, _vp]
    lib.fa_b200_exchange_combine.restype = C.c_int
    lib.fa_b200_exchange_combine.argtypes = [C.POINTER(FaExchangeParams), _vp]
    lib.fa_b200_exchange_params_size.restype = C.c_size_t
    if lib.fa_b200_exchange_params_size() != C.sizeof(FaExchangeParams):
        raise ImportError("FaExchangeParams layout mismatch between libfa_b200.so and its binding")
    lib.fa_b200_fused_exchange_size.restype = C.c_size_t
    if lib.fa_b200_fused_exchange_size() != C.sizeof(FaFusedExchange):
        raise ImportError("FaFusedExchange layout mismatch between libfa_b200.so and its binding")
    if lib.fa_b200_abi_version() != ABI_VERSION:
        raise ImportError("libfa_b200.so ABI version mismatch")
    if lib.fa_b200_params_size() != C.sizeof(FaFwdParams):
        raise ImportError(
            f"FaFwdParams layout mismatch: library {lib.fa_b200_params_size()} B, binding {C.sizeof(FaFwdParams)} B")
    return lib


lib = _load()


def last_error() -> str:
    return lib.fa_b200_last_error().decode("utf-8", "replace")


def check(rc: int) -> None:
    """Turn a negative return code into RuntimeError — the reference's TORCH_CHECK behaviour
    (flash_api.cpp:16-18 surfaces as RuntimeError in Python)."""
    if rc != FA_OK:
        raise RuntimeError(last_error() or f"libfa_b200 error {rc}")
