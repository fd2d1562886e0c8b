"""Shared helpers for the test-suite (TEST INFRASTRUCTURE)."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "vllm-flash-attention_b200"), os.path.join(ROOT, "oracle"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libfa2_ref_sm100.so")
_ref = None


def ref_lib():
    """The reference's own FA2 kernels built for sm_100 (oracle/Makefile), or None."""
    global _ref
    if _ref is None and os.path.exists(REF_LIB):
        import torch  # noqa: F401  (libc10 must be loaded first)
        lib = C.CDLL(REF_LIB)
        lib.ref_fa2_last_error.restype = C.c_char_p
        lib.ref_fa2_workspace_bytes.restype = C.c_size_t
        lib.ref_fa2_run.restype = C.c_int
        _ref = lib
    return _ref


def ref_kernel_fwd(q, k, v, causal=False, window=(-1, -1), alibi_slopes=None, softmax_scale=None):
    """Reference FA2 kernel, dense (b,s,h,d) — mha_fwd semantics."""
    from vllm_flash_attn._cabi import FaFwdParams
    lib = ref_lib()
    b, sq, h, d = q.shape
    sk, hk = k.shape[1], k.shape[2]
    out = torch.empty_like(q)
    lse = torch.empty(b, h, sq, dtype=torch.float32, device=q.device)
    P = FaFwdParams()
    P.q, P.k, P.v, P.o, P.softmax_lse = q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), lse.data_ptr()
    P.q_batch_stride, P.q_row_stride, P.q_head_stride = q.stride(0), q.stride(1), q.stride(2)
    P.k_batch_stride, P.k_row_stride, P.k_head_stride = k.stride(0), k.stride(1), k.stride(2)
    P.v_batch_stride, P.v_row_stride, P.v_head_stride = v.stride(0), v.stride(1), v.stride(2)
    P.o_batch_stride, P.o_row_stride, P.o_head_stride = out.stride(0), out.stride(1), out.stride(2)
    P.b, P.seqlen_q, P.seqlen_k, P.h, P.h_k, P.d = b, sq, sk, h, hk, d
    P.is_bf16 = int(q.dtype == torch.bfloat16)
    P.softmax_scale = softmax_scale if softmax_scale is not None else d ** -0.5
    P.is_causal, P.window_left, P.window_right = int(causal), window[0], window[1]
    P.is_seqlens_k_cumulative = 1
    if alibi_slopes is not None:
        P.alibi_slopes = alibi_slopes.data_ptr()
        P.alibi_slopes_batch_stride = alibi_slopes.stride(0) if alibi_slopes.dim() == 2 else 0
    need = lib.ref_fa2_workspace_bytes(C.byref(P))
    ws = torch.empty(need, dtype=torch.uint8, device=q.device)
    P.workspace, P.workspace_bytes = ws.data_ptr(), need
    rc = lib.ref_fa2_run(C.byref(P), 0, C.c_void_p(torch.cuda.current_stream().cuda_stream))
    if rc != 0:
        raise RuntimeError(lib.ref_fa2_last_error().decode())
    return out, lse


def ref_kernel_varlen(q, k, v, cu_q, cu_k, max_q, max_k, causal=False, window=(-1, -1), block_table=None,
                      softmax_scale=None):
    """Reference FA2 kernel, varlen (+ paged) — mha_varlen_fwd semantics."""
    from vllm_flash_attn._cabi import FaFwdParams
    lib = ref_lib()
    total_q, h, d = q.shape
    b = cu_q.numel() - 1
    out = torch.empty_like(q)
    lse = torch.empty(b, h, max_q, dtype=torch.float32, device=q.device)
    P = FaFwdParams()
    P.q, P.k, P.v, P.o, P.softmax_lse = q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), lse.data_ptr()
    P.q_row_stride, P.q_head_stride = q.stride(0), q.stride(1)
    P.o_row_stride, P.o_head_stride = out.stride(0), out.stride(1)
    if block_table is not None:
        hk = k.shape[2]
        P.k_batch_stride, P.k_row_stride, P.k_head_stride = k.stride(0), k.stride(1), k.stride(2)
        P.v_batch_stride, P.v_row_stride, P.v_head_stride = v.stride(0), v.stride(1), v.stride(2)
        P.block_table, P.block_table_batch_stride = block_table.data_ptr(), block_table.stride(0)
        P.page_block_size = k.shape[1]
    else:
        hk = k.shape[1]
        P.k_row_stride, P.k_head_stride = k.stride(0), k.stride(1)
        P.v_row_stride, P.v_head_stride = v.stride(0), v.stride(1)
    P.b, P.seqlen_q, P.seqlen_k, P.h, P.h_k, P.d = b, max_q, max_k, h, hk, d
    P.is_bf16 = int(q.dtype == torch.bfloat16)
    P.softmax_scale = softmax_scale if softmax_scale is not None else d ** -0.5
    P.is_causal, P.window_left, P.window_right = int(causal), window[0], window[1]
    P.cu_seqlens_q, P.cu_seqlens_k = cu_q.data_ptr(), cu_k.data_ptr()
    P.is_seqlens_k_cumulative = 1
    # mha_varlen_fwd splits only in its GQA-swapped pure-decode form (flash_api.cpp:618-623); the glue's size query always
    # reports the 128-split worst case, which for a long prefill batch is terabytes nobody will touch
    need = lib.ref_fa2_workspace_bytes(C.byref(P)) if max_q == 1 else 16
    ws = torch.empty(need, dtype=torch.uint8, device=q.device)
    P.workspace, P.workspace_bytes = ws.data_ptr(), need
    rc = lib.ref_fa2_run(C.byref(P), 1, C.c_void_p(torch.cuda.current_stream().cuda_stream))
    if rc != 0:
        raise RuntimeError(lib.ref_fa2_last_error().decode())
    return out, lse


def ref_kernel_kvcache(q, kc, vc, cache_seqlens, block_table=None, causal=False, window=(-1, -1), num_splits=0,
                       softmax_scale=None):
    """Reference FA2 kernel — mha_fwd_kvcache semantics (no append)."""
    from vllm_flash_attn._cabi import FaFwdParams
    lib = ref_lib()
    b, sq, h, d = q.shape
    hk = kc.shape[2]
    out = torch.empty_like(q)
    lse = torch.empty(b, h, sq, dtype=torch.float32, device=q.device)
    P = FaFwdParams()
    P.q, P.k, P.v, P.o, P.softmax_lse = q.data_ptr(), kc.data_ptr(), vc.data_ptr(), out.data_ptr(), lse.data_ptr()
    P.q_batch_stride, P.q_row_stride, P.q_head_stride = q.stride(0), q.stride(1), q.stride(2)
    P.k_batch_stride, P.k_row_stride, P.k_head_stride = kc.stride(0), kc.stride(1), kc.stride(2)
    P.v_batch_stride, P.v_row_stride, P.v_head_stride = vc.stride(0), vc.stride(1), vc.stride(2)
    P.o_batch_stride, P.o_row_stride, P.o_head_stride = out.stride(0), out.stride(1), out.stride(2)
    sk = kc.shape[1] if block_table is None else block_table.shape[1] * kc.shape[1]
    P.b, P.seqlen_q, P.seqlen_k, P.h, P.h_k, P.d = b, sq, sk, h, hk, d
    P.is_bf16 = int(q.dtype == torch.bfloat16)
    P.softmax_scale = softmax_scale if softmax_scale is not None else d ** -0.5
    P.is_causal, P.window_left, P.window_right = int(causal), window[0], window[1]
    P.cu_seqlens_k = cache_seqlens.data_ptr()
    P.is_seqlens_k_cumulative = 0
    if block_table is not None:
        P.block_table, P.block_table_batch_stride = block_table.data_ptr(), block_table.stride(0)
        P.page_block_size = kc.shape[1]
    P.num_splits = num_splits
    need = lib.ref_fa2_workspace_bytes(C.byref(P))
    ws = torch.empty(need, dtype=torch.uint8, device=q.device)
    P.workspace, P.workspace_bytes = ws.data_ptr(), need
    rc = lib.ref_fa2_run(C.byref(P), 2, C.c_void_p(torch.cuda.current_stream().cuda_stream))
    if rc != 0:
        raise RuntimeError(lib.ref_fa2_last_error().decode())
    return out, lse
