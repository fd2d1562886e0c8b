"""Drop-in for the reference's pybind11 extension ``vllm_flash_attn_2_cuda``
(csrc/flash_attn/flash_api.cpp:939-946): the same three ops with the same positional arguments
and return lists, implemented as a thin tensor-unwrapping shim over the C ABI of libfa_b200.so.

What lives here is only what needs torch: device/dtype/shape checks on tensors, allocation of
``out`` / ``softmax_lse`` / split workspace, and reading strides.  Everything else (mask
normalisation, kernel choice, split heuristic, launches) is in the library (csrc/fa_api.cu).
Kernels are enqueued on torch's current CUDA stream of q's device, no host synchronisation
(flash_api.cpp:344,399).

Two host bindings over the same library: when the compiled pybind11 / ATen module ``_fa_torch`` (csrc/fa_torch_ext.cpp,
built by build_torch_ext.sh — the form the reference's own extension has) is present, the three plain ops go through it
(a few microseconds of host time per call instead of ~25 through ctypes); the Python-over-ctypes implementation below
is the same logic, serves when that module was not built, and carries the sharding / test extensions
(``_partial_buf``, ``_kv_shard``, ``_fused_exchange``, ``_force_kernel``).  ``FA_B200_HOST=ctypes`` forces it.
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from typing import List, Optional

import torch

from . import _cabi
from ._cabi import FaFwdParams, OP_FWD, OP_KVCACHE, OP_VARLEN, check, lib

_force_kernel = 0          # tests: _cabi.FA_KERNEL_* to pin a kernel family
_two_step_append = False   # tests: always run the stand-alone append kernel (ctypes path only)

host_binding_note = ""      # why the compiled binding is not in use, if it is not
try:
    from . import _fa_torch as _N        # compiled host binding (optional; same library underneath)
    if _N.abi_version() != _cabi.ABI_VERSION:
        host_binding_note = (f"_fa_torch was built against ABI {_N.abi_version()}, libfa_b200.so has {_cabi.ABI_VERSION}: "
                             "rebuild with build_torch_ext.sh")
        _N = None
except (ImportError, OSError) as e:      # not built, or built against another torch: the ctypes binding serves
    _N = None
    host_binding_note = f"{type(e).__name__}: {e}"
_NATIVE = _N               # the loaded module, whatever _N is switched to later (tests toggle _N)
if os.environ.get("FA_B200_HOST", "") == "ctypes":
    _N = None
host_binding = "pybind11" if _N is not None else "ctypes"


class _PerThreadInfo:
    """Mapping with one dict per host thread: the diagnostics of the most recent call ON THIS THREAD (kernel_used,
    num_splits, launches).  The reference's extension keeps no global mutable state (flash_api.cpp: plain functions),
    so neither does this module: two threads driving two streams never see each other's values."""

    def __init__(self):
        self._tls = threading.local()

    def _d(self):
        d = getattr(self._tls, "d", None)
        if d is None:
            d = self._tls.d = {}
        if getattr(self._tls, "native", False):      # the last call on this thread went through _fa_torch
            self._tls.native = False
            ku, ns, ln = _NATIVE.last_call_info()
            d.update(kernel_used=ku, num_splits=ns, launches=ln)
        return d

    def _native_call(self):
        self._tls.native = True

    def __getitem__(self, k):
        return self._d()[k]

    def get(self, k, default=None):
        return self._d().get(k, default)

    def update(self, *a, **kw):
        self._d().update(*a, **kw)

    def keys(self):
        return self._d().keys()

    def items(self):
        return self._d().items()

    def __iter__(self):
        return iter(self._d())

    def __len__(self):
        return len(self._d())

    def __repr__(self):
        return repr(self._d())


last_call_info = _PerThreadInfo()


def _chk(cond: bool, msg: str) -> None:
    if not cond:
        raise RuntimeError(msg)


def _check_qkv(q, k, v):
    _chk(q.dtype in (torch.float16, torch.bfloat16), "FlashAttention only support fp16 and bf16 data type")
    _chk(k.dtype == q.dtype, "query and key must have the same dtype")
    _chk(v.dtype == q.dtype, "query and value must have the same dtype")
    for t, n in ((q, "q"), (k, "k"), (v, "v")):
        _chk(t.is_cuda, f"{n} must be on CUDA")
        _chk(t.stride(-1) == 1, "Input tensor must have contiguous last dimension")


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _stream(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def _i32_dev(t, name, device):
    _chk(t.dtype == torch.int32, f"{name} must have dtype int32")
    _chk(t.is_cuda and t.device == device, f"{name} must be on the same CUDA device as q")
    return t


def _alibi(P, alibi_slopes, batch, heads, device):
    if alibi_slopes is None:
        return
    _chk(alibi_slopes.dtype == torch.float32, "ALiBi slopes must have dtype fp32")
    _chk(alibi_slopes.is_cuda and alibi_slopes.device == device, "ALiBi slopes must be on the same device")
    _chk(alibi_slopes.stride(-1) == 1, "ALiBi slopes tensor must have contiguous last dimension")
    _chk(tuple(alibi_slopes.shape) in ((heads,), (batch, heads)), "ALiBi slopes must have shape (nheads,) or (batch, nheads)")
    P.alibi_slopes = alibi_slopes.data_ptr()
    P.alibi_slopes_batch_stride = alibi_slopes.stride(0) if alibi_slopes.dim() == 2 else 0


_ws_cache = {}             # (device index, stream handle, host thread) -> persistent workspace tensor (zero-initialised)


def _run(P: FaFwdParams, op: int, fn, device) -> None:
    """One library call.  The split-KV / work-list workspace is a persistent per-(device, stream, host thread)
    buffer (calls one thread makes on one stream are ordered, so they can share it; two threads issuing on the
    same stream get a buffer each and never race on the prefill work list), grown when the library says it is
    too small; under CUDA-graph capture a private buffer sized by fa_b200_workspace_bytes is used
    instead, so that the capture owns it."""
    P.force_kernel = _force_kernel
    stream = _stream(device)
    ref = C.byref(P)
    if torch.cuda.is_current_stream_capturing():
        need = lib.fa_b200_workspace_bytes(ref, op)
        if need == C.c_size_t(-1).value:
            check(-1)
        ws = torch.zeros(max(need, 16), dtype=torch.uint8, device=device)     # zeroed: workspace_clean (a memset node of the capture)
        P.workspace, P.workspace_bytes, P.workspace_clean = ws.data_ptr(), ws.numel(), 1
        check(fn(ref, stream))
    else:
        key = (device.index, stream, threading.get_ident())
        ws = _ws_cache.get(key)
        P.workspace_clean = 1       # the cached buffers are allocated zeroed and only the library writes to them
        if ws is not None:
            P.workspace, P.workspace_bytes = ws.data_ptr(), ws.numel()
        rc = fn(ref, stream)
        if rc == _cabi.FA_ERR_WORKSPACE:
            need = lib.fa_b200_workspace_bytes(ref, op)
            if need == C.c_size_t(-1).value:
                check(-1)
            ws = torch.zeros(max(2 * need, 1 << 20), dtype=torch.uint8, device=device)
            _ws_cache[key] = ws
            P.workspace, P.workspace_bytes = ws.data_ptr(), ws.numel()
            rc = fn(ref, stream)
        check(rc)
    last_call_info.update(kernel_used=P.kernel_used, num_splits=P.num_splits, launches=P.launches)


class _on_device:
    """`with torch.cuda.device(dev)` only when dev is not already current (the guard costs microseconds)."""
    __slots__ = ("g",)

    def __init__(self, dev):
        self.g = None if torch.cuda.current_device() == dev.index else torch.cuda.device(dev)

    def __enter__(self):
        if self.g is not None:
            self.g.__enter__()

    def __exit__(self, *a):
        if self.g is not None:
            self.g.__exit__(*a)


_placeholder_cache = {}


def _placeholders(dtype, device):
    """The `p` (empty: no dropout, so no returned probabilities) and `rng_state` outputs of fwd / varlen_fwd
    (flash_api.cpp:384-392,409).  Nothing ever writes them, so one pair per (dtype, device) is shared by
    all calls instead of two allocator round trips per call."""
    key = (dtype, device)
    t = _placeholder_cache.get(key)
    if t is None:
        t = (torch.empty((0,), dtype=dtype, device=device), torch.zeros((2,), dtype=torch.int64, device=device))
        _placeholder_cache[key] = t
    return t


def _pad_head(x, d_pad):
    return torch.nn.functional.pad(x, (0, d_pad - x.shape[-1])) if x.shape[-1] != d_pad else x


def fwd(q, k, v, out_, alibi_slopes_, p_dropout, softmax_scale, is_causal, window_size_left,
        window_size_right, return_softmax, gen_) -> List[torch.Tensor]:
    """mha_fwd (flash_api.cpp:240-420): compiled binding when present, else the ctypes implementation below."""
    if _N is not None and _force_kernel == 0:
        r = _N.fwd(q, k, v, out_, alibi_slopes_, p_dropout, softmax_scale, is_causal, window_size_left,
                   window_size_right, return_softmax, gen_)
        last_call_info._native_call()
        return r
    return _fwd_py(q, k, v, out_, alibi_slopes_, p_dropout, softmax_scale, is_causal, window_size_left,
                   window_size_right, return_softmax, gen_)


def _fwd_py(q, k, v, out_, alibi_slopes_, p_dropout, softmax_scale, is_causal, window_size_left,
            window_size_right, return_softmax, gen_) -> List[torch.Tensor]:
    """mha_fwd (flash_api.cpp:240-420).  q (b,sq,h,d); k,v (b,sk,hk,d).
    Returns [out, q_padded, k_padded, v_padded, out_padded, softmax_lse, p, rng_state]."""
    _check_qkv(q, k, v)
    _chk(p_dropout == 0.0, "This flash attention build does not support dropout.")
    _chk(q.dim() == 4 and k.dim() == 4 and v.dim() == 4, "q, k, v must be (batch, seqlen, nheads, headdim)")
    b, sq, h, d_og = q.shape
    sk, hk = k.shape[1], k.shape[2]
    _chk(b > 0, "batch size must be postive")
    _chk(d_og <= 256, "FlashAttention forward only supports head dimension at most 256")
    _chk(h % hk == 0, "Number of heads in key/value must divide number of heads in query")
    _chk(tuple(k.shape) == (b, sk, hk, d_og) and tuple(v.shape) == (b, sk, hk, d_og), "k/v shape mismatch")
    d = (d_og + 7) // 8 * 8
    q_p, k_p, v_p = _pad_head(q, d), _pad_head(k, d), _pad_head(v, d)
    if out_ is not None:
        _chk(out_.dtype == q.dtype, "Output must have the same dtype as inputs")
        _chk(out_.is_cuda and out_.device == q.device, "Output must be on the same device")
        _chk(out_.stride(-1) == 1, "Output tensor must have contiguous last dimension")
        _chk(tuple(out_.shape) == (b, sq, h, d_og), "out has the wrong shape")
        out = out_ if d == d_og else torch.empty_like(q_p)
    else:
        out = torch.empty_like(q_p)
    lse = torch.empty((b, h, sq), dtype=torch.float32, device=q.device)
    with _on_device(q.device):
        if sk > 0 and sq > 0:
            P = FaFwdParams()
            P.q, P.k, P.v, P.o, P.softmax_lse = q_p.data_ptr(), k_p.data_ptr(), v_p.data_ptr(), out.data_ptr(), lse.data_ptr()
            P.q_batch_stride, P.q_row_stride, P.q_head_stride = q_p.stride(0), q_p.stride(1), q_p.stride(2)
            P.k_batch_stride, P.k_row_stride, P.k_head_stride = k_p.stride(0), k_p.stride(1), k_p.stride(2)
            P.v_batch_stride, P.v_row_stride, P.v_head_stride = v_p.stride(0), v_p.stride(1), v_p.stride(2)
            P.o_batch_stride, P.o_row_stride, P.o_head_stride = out.stride(0), out.stride(1), out.stride(2)
            P.b, P.seqlen_q, P.seqlen_k, P.h, P.h_k, P.d = b, sq, sk, h, hk, d
            P.is_bf16 = int(q.dtype == torch.bfloat16)
            P.softmax_scale = softmax_scale
            P.is_causal, P.window_left, P.window_right = int(is_causal), window_size_left, window_size_right
            P.is_seqlens_k_cumulative = 1
            _alibi(P, alibi_slopes_, b, h, q.device)
            _run(P, OP_FWD, lib.fa_b200_fwd, q.device)
        else:
            out.zero_()                       # flash_api.cpp:401-405
            lse.fill_(float("inf"))
    out_padded = out
    if d != d_og:
        out = out[..., :d_og]
        if out_ is not None:
            out_.copy_(out)
            out = out_
    p, rng_state = _placeholders(q.dtype, q.device)                 # return_softmax needs dropout (off)
    return [out, q_p, k_p, v_p, out_padded, lse, p, rng_state]


def varlen_fwd(q, k, v, out_, cu_seqlens_q, cu_seqlens_k, seqused_k, block_table_, alibi_slopes_,
               max_seqlen_q, max_seqlen_k, p_dropout, softmax_scale, zero_tensors, is_causal,
               window_size_left, window_size_right, return_softmax, gen_) -> List[torch.Tensor]:
    """mha_varlen_fwd (flash_api.cpp:422-669): compiled binding when present, else the ctypes implementation below."""
    if _N is not None and _force_kernel == 0:
        r = _N.varlen_fwd(q, k, v, out_, cu_seqlens_q, cu_seqlens_k, seqused_k, block_table_, alibi_slopes_,
                          max_seqlen_q, max_seqlen_k, p_dropout, softmax_scale, zero_tensors, is_causal,
                          window_size_left, window_size_right, return_softmax, gen_)
        last_call_info._native_call()
        return r
    return _varlen_fwd_py(q, k, v, out_, cu_seqlens_q, cu_seqlens_k, seqused_k, block_table_, alibi_slopes_,
                          max_seqlen_q, max_seqlen_k, p_dropout, softmax_scale, zero_tensors, is_causal,
                          window_size_left, window_size_right, return_softmax, gen_)


def _varlen_fwd_py(q, k, v, out_, cu_seqlens_q, cu_seqlens_k, seqused_k, block_table_, alibi_slopes_,
                   max_seqlen_q, max_seqlen_k, p_dropout, softmax_scale, zero_tensors, is_causal,
                   window_size_left, window_size_right, return_softmax, gen_) -> List[torch.Tensor]:
    """mha_varlen_fwd (flash_api.cpp:422-669).  q (total_q,h,d); k,v (total_k,hk,d) or, with
    block_table (b,max_blocks), paged (num_blocks,page,hk,d).  softmax_lse is (b,h,max_seqlen_q)."""
    _check_qkv(q, k, v)
    _chk(p_dropout == 0.0, "This flash attention build does not support dropout.")
    dev = q.device
    _i32_dev(cu_seqlens_q, "cu_seqlens_q", dev); _i32_dev(cu_seqlens_k, "cu_seqlens_k", dev)
    _chk(cu_seqlens_q.is_contiguous() and cu_seqlens_k.is_contiguous(), "cu_seqlens must be contiguous")
    paged = block_table_ is not None
    _chk(q.dim() == 3, "q must be (total_q, nheads, headdim)")
    total_q, h, d_og = q.shape
    b = cu_seqlens_q.numel() - 1
    _chk(b > 0, "batch size must be positive")
    _chk(cu_seqlens_k.numel() == b + 1, "cu_seqlens_k must have shape (batch_size + 1)")
    _chk(d_og <= 256, "FlashAttention forward only supports head dimension at most 256")
    if paged:
        _i32_dev(block_table_, "block_table", dev)
        _chk(block_table_.stride(-1) == 1, "block_table must have contiguous last dimension")
        _chk(k.dim() == 4 and v.dim() == 4, "paged k/v must be (num_blocks, page_block_size, nheads_k, headdim)")
        num_blocks, page, hk = k.shape[0], k.shape[1], k.shape[2]
        _chk(page % 16 == 0, "Paged KV cache block size must be divisible by 16")
        _chk(tuple(block_table_.shape[:1]) == (b,), "block_table must have shape (batch_size, max_num_blocks_per_seq)")
        _chk(tuple(v.shape) == tuple(k.shape) and k.shape[3] == d_og, "k/v shape mismatch")
    else:
        _chk(k.dim() == 3 and v.dim() == 3, "k, v must be (total_k, nheads_k, headdim)")
        hk = k.shape[1]
        _chk(tuple(v.shape) == tuple(k.shape) and k.shape[2] == d_og, "k/v shape mismatch")
    _chk(h % hk == 0, "Number of heads in key/value must divide number of heads in query")
    if seqused_k is not None:
        _i32_dev(seqused_k, "seqused_k", dev)
        _chk(seqused_k.is_contiguous() and seqused_k.numel() == b, "seqused_k must have shape (batch_size)")
    d = (d_og + 7) // 8 * 8
    q_p, k_p, v_p = _pad_head(q, d), _pad_head(k, d), _pad_head(v, d)
    if out_ is not None:
        _chk(out_.dtype == q.dtype, "Output must have the same dtype as inputs")
        _chk(out_.is_cuda and out_.device == dev, "Output must be on the same device")
        _chk(out_.stride(-1) == 1, "Output tensor must have contiguous last dimension")
        _chk(tuple(out_.shape) == (total_q, h, d_og), "out has the wrong shape")
        out = out_ if d == d_og else torch.empty_like(q_p)
    else:
        out = torch.empty_like(q_p)
    lse = torch.empty((b, h, max_seqlen_q), dtype=torch.float32, device=dev)
    if zero_tensors:
        out.zero_(); lse.fill_(float("-inf"))
    with _on_device(dev):
        if max_seqlen_k > 0 and max_seqlen_q > 0 and total_q > 0:
            P = FaFwdParams()
            P.q, P.k, P.v, P.o, P.softmax_lse = q_p.data_ptr(), k_p.data_ptr(), v_p.data_ptr(), out.data_ptr(), lse.data_ptr()
            P.q_row_stride, P.q_head_stride = q_p.stride(0), q_p.stride(1)
            P.o_row_stride, P.o_head_stride = out.stride(0), out.stride(1)
            if paged:
                P.k_batch_stride, P.k_row_stride, P.k_head_stride = k_p.stride(0), k_p.stride(1), k_p.stride(2)
                P.v_batch_stride, P.v_row_stride, P.v_head_stride = v_p.stride(0), v_p.stride(1), v_p.stride(2)
                P.block_table = block_table_.data_ptr()
                P.block_table_batch_stride = block_table_.stride(0)
                P.page_block_size, P.max_blocks_per_seq, P.num_blocks = page, block_table_.shape[1], num_blocks
                P.total_k = 0
            else:
                P.k_row_stride, P.k_head_stride = k_p.stride(0), k_p.stride(1)
                P.v_row_stride, P.v_head_stride = v_p.stride(0), v_p.stride(1)
                P.total_k = k_p.shape[0]
            P.total_q = total_q
            P.b, P.seqlen_q, P.seqlen_k, P.h, P.h_k, P.d = b, max_seqlen_q, max_seqlen_k, h, hk, d
            P.is_bf16 = int(q.dtype == torch.bfloat16)
            P.softmax_scale = softmax_scale
            P.is_causal, P.window_left, P.window_right = int(is_causal), window_size_left, window_size_right
            P.cu_seqlens_q, P.cu_seqlens_k = cu_seqlens_q.data_ptr(), cu_seqlens_k.data_ptr()
            P.seqused_k = _ptr(seqused_k)
            P.is_seqlens_k_cumulative = 1
            _alibi(P, alibi_slopes_, b, h, dev)
            _run(P, OP_VARLEN, lib.fa_b200_varlen_fwd, dev)
        else:
            out.zero_(); lse.fill_(float("inf"))
    out_padded = out
    if d != d_og:
        out = out[..., :d_og]
        if out_ is not None:
            out_.copy_(out)
            out = out_
    p, rng_state = _placeholders(q.dtype, dev)
    return [out, q_p, k_p, v_p, out_padded, lse, p, rng_state]


def fwd_kvcache(q, kcache, vcache, k_, v_, seqlens_k_, rotary_cos_, rotary_sin_, cache_batch_idx_,
                block_table_, alibi_slopes_, out_, softmax_scale, is_causal, window_size_left,
                window_size_right, is_rotary_interleaved, num_splits, *, _partial_buf=None, _kv_shard=None,
                _fused_exchange=None) -> List[torch.Tensor]:
    """mha_fwd_kvcache (flash_api.cpp:671-937).  q (b,sq,h,d); cache (b_c,sk,hk,d) or paged
    (num_blocks,page,hk,d) with block_table (b,max_blocks); optional new k/v (b,sk_new,hk,d) appended
    in place at cache_seqlens (with rotary on k and q).  Returns [out, softmax_lse].

    ``_partial_buf`` (extension used by vllm_flash_attn.sharded): a 1-D fp32 tensor of
    b*h*sq*(d+1) elements; the call then runs in the C ABI's partial_out mode and writes the packed
    record [o (b,h,sq,d) | lse (b,h,sq)] into it (lse = -inf for rows with no key in this shard),
    returning views of the two parts.

    ``_fused_exchange`` (extension used by vllm_flash_attn.sharded): a _cabi.FaFusedExchange; with ``_kv_shard`` set the
    decode kernel then merges its splits, exchanges the merged rows with the peer GPUs and writes the FINAL out / lse
    itself (one launch per decode step)."""
    if (_N is not None and _force_kernel == 0 and not _two_step_append and _partial_buf is None and _kv_shard is None
            and _fused_exchange is None):
        r = _N.fwd_kvcache(q, kcache, vcache, k_, v_, seqlens_k_, rotary_cos_, rotary_sin_, cache_batch_idx_, block_table_,
                           alibi_slopes_, out_, softmax_scale, is_causal, window_size_left, window_size_right,
                           is_rotary_interleaved, num_splits)
        last_call_info._native_call()
        return r
    _check_qkv(q, kcache, vcache)
    dev = q.device
    _chk(q.dim() == 4 and kcache.dim() == 4 and vcache.dim() == 4, "q, kcache, vcache must be 4-D")
    b, sq, h, d_og = q.shape
    paged = block_table_ is not None
    if paged:
        _chk(cache_batch_idx_ is None, "Paged KVcache does not support cache_batch_idx")
        _i32_dev(block_table_, "block_table", dev)
        _chk(block_table_.stride(-1) == 1, "block_table must have contiguous last dimension")
        num_blocks, page, hk = kcache.shape[0], kcache.shape[1], kcache.shape[2]
        _chk(page % 16 == 0, "Paged KV cache block size must be divisible by 16")
        _chk(block_table_.shape[0] == b, "block_table must have shape (batch_size, max_num_blocks_per_seq)")
        max_blocks = block_table_.shape[1]
        sk = max_blocks * page
    else:
        b_c, sk, hk = kcache.shape[0], kcache.shape[1], kcache.shape[2]
        num_blocks = page = max_blocks = 0
        if cache_batch_idx_ is None:
            _chk(b_c == b, "kcache batch size must match q (or pass cache_batch_idx)")
    _chk(b > 0, "batch size must be postive")
    _chk(d_og <= 256, "FlashAttention forward only supports head dimension at most 256")
    _chk(h % hk == 0, "Number of heads in key/value must divide number of heads in query")
    _chk(tuple(vcache.shape) == tuple(kcache.shape) and kcache.shape[3] == d_og, "kcache/vcache shape mismatch")
    _chk(d_og % 8 == 0, "fwd_kvcache: head dimension must be a multiple of 8 (cache cannot be padded in place)")
    d = d_og
    partial = _partial_buf is not None
    if partial:
        rows = b * h * sq
        _chk(out_ is None, "out= cannot be combined with partial mode")
        _chk(_partial_buf.dtype == torch.float32 and _partial_buf.is_cuda and _partial_buf.device == dev and
             _partial_buf.is_contiguous() and _partial_buf.numel() == rows * (d + 1),
             "partial buffer must be a contiguous fp32 CUDA tensor of b*h*sq*(d+1) elements")
        out = _partial_buf[: rows * d].view(b, h, sq, d)
        lse = _partial_buf[rows * d:].view(b, h, sq)
    elif out_ is not None:
        _chk(out_.dtype == q.dtype, "Output must have the same dtype as inputs")
        _chk(out_.is_cuda and out_.device == dev, "Output must be on the same device")
        _chk(out_.stride(-1) == 1, "Output tensor must have contiguous last dimension")
        # The reference checks `out` against the GQA-swapped shape (flash_api.cpp:787), which rejects
        # a normal (b,sq,h,d) out whenever its packing triggers; we accept the natural shape.
        _chk(tuple(out_.shape) == (b, sq, h, d), "out has the wrong shape")
        out = out_
    else:
        out = torch.empty_like(q)
    if not partial:
        lse = torch.empty((b, h, sq), dtype=torch.float32, device=dev)
    if seqlens_k_ is not None:
        _i32_dev(seqlens_k_, "seqlens_k", dev)
        _chk(seqlens_k_.is_contiguous() and tuple(seqlens_k_.shape) == (b,), "seqlens_k must have shape (batch_size)")
    if cache_batch_idx_ is not None:
        _i32_dev(cache_batch_idx_, "cache_batch_idx", dev)
        _chk(cache_batch_idx_.is_contiguous() and tuple(cache_batch_idx_.shape) == (b,), "cache_batch_idx must have shape (batch_size)")

    with _on_device(dev):
        P = FaFwdParams()
        P.q, P.k, P.v, P.o, P.softmax_lse = q.data_ptr(), kcache.data_ptr(), vcache.data_ptr(), out.data_ptr(), lse.data_ptr()
        P.q_batch_stride, P.q_row_stride, P.q_head_stride = q.stride(0), q.stride(1), q.stride(2)
        P.k_batch_stride, P.k_row_stride, P.k_head_stride = kcache.stride(0), kcache.stride(1), kcache.stride(2)
        P.v_batch_stride, P.v_row_stride, P.v_head_stride = vcache.stride(0), vcache.stride(1), vcache.stride(2)
        if partial:   # (b,h,sq,d) fp32
            P.o_batch_stride, P.o_row_stride, P.o_head_stride = out.stride(0), out.stride(2), out.stride(1)
            P.partial_out = 1
        else:
            P.o_batch_stride, P.o_row_stride, P.o_head_stride = out.stride(0), out.stride(1), out.stride(2)
        if _kv_shard is not None:           # (start, end): the cache holds that key range, seqlens_k are global
            P.kv_shard_start, P.kv_shard_len = int(_kv_shard[0]), int(_kv_shard[1] - _kv_shard[0])
        if _fused_exchange is not None:
            P.fused_exchange = C.addressof(_fused_exchange)
        P.b, P.seqlen_q, P.seqlen_k, P.h, P.h_k, P.d = b, sq, sk, h, hk, d
        P.is_bf16 = int(q.dtype == torch.bfloat16)
        P.softmax_scale = softmax_scale
        P.is_causal, P.window_left, P.window_right = int(is_causal), window_size_left, window_size_right
        P.cu_seqlens_k = _ptr(seqlens_k_)
        P.is_seqlens_k_cumulative = 0 if seqlens_k_ is not None else 1
        P.cache_batch_idx = _ptr(cache_batch_idx_)
        if paged:
            P.block_table, P.block_table_batch_stride = block_table_.data_ptr(), block_table_.stride(0)
            P.page_block_size, P.max_blocks_per_seq, P.num_blocks = page, max_blocks, num_blocks
        else:
            P.num_blocks = b_c          # batch entries of the dense cache (extent of the K/V tensor maps)
        _alibi(P, alibi_slopes_, b, h, dev)
        P.num_splits = num_splits

        q_rot = None
        if k_ is not None:
            # ---- append new K/V into the cache (+ rotary on k and q), flash_fwd_kernel.h:672-824 ----
            _chk(v_ is not None, "key and value must be passed together")
            _chk(seqlens_k_ is not None, "seqlens_k must be passed if key/value are passed")
            _chk(k_.dtype == q.dtype and v_.dtype == q.dtype, "Key and value must have the same dtype as query")
            _chk(k_.is_cuda and v_.is_cuda and k_.stride(-1) == 1 and v_.stride(-1) == 1, "new k/v must be CUDA, last dim contiguous")
            sk_new = k_.shape[1]
            _chk(tuple(k_.shape) == (b, sk_new, hk, d) and tuple(v_.shape) == (b, sk_new, hk, d), "new k/v shape mismatch")
            P.knew, P.vnew, P.seqlen_knew = k_.data_ptr(), v_.data_ptr(), sk_new
            P.knew_batch_stride, P.knew_row_stride, P.knew_head_stride = k_.stride(0), k_.stride(1), k_.stride(2)
            P.vnew_batch_stride, P.vnew_row_stride, P.vnew_head_stride = v_.stride(0), v_.stride(1), v_.stride(2)
            if rotary_cos_ is not None:
                _chk(rotary_sin_ is not None, "rotary_cos and rotary_sin must be passed together")
                _chk(rotary_cos_.dtype == q.dtype and rotary_sin_.dtype == q.dtype, "rotary_cos/sin must have the same dtype as query")
                _chk(rotary_cos_.is_contiguous() and rotary_sin_.is_contiguous(), "rotary_cos/sin must be contiguous")
                rd = rotary_cos_.shape[1] * 2
                _chk(rd <= d and rd % 16 == 0, "rotary_dim must be <= headdim and a multiple of 16")
                _chk(rotary_cos_.shape[0] >= sk, "cos/sin seqlen must be at least the seqlen of KV cache")
                _chk(tuple(rotary_sin_.shape) == tuple(rotary_cos_.shape), "rotary_sin shape mismatch")
                P.rotary_cos, P.rotary_sin = rotary_cos_.data_ptr(), rotary_sin_.data_ptr()
                P.rotary_dim, P.seqlen_ro = rd, rotary_cos_.shape[0]
                P.is_rotary_interleaved = int(is_rotary_interleaved)
            P.force_kernel = _force_kernel
            fused = lib.fa_b200_fused_append_supported(C.byref(P)) if not _two_step_append else 0
            if fused < 0:
                check(fused)
            if not fused:
                # not a decode shape (or a sliding window): the stand-alone append kernel first, then attention over
                # the grown cache with the rotated q
                if rotary_cos_ is not None:
                    q_rot = torch.empty((b, sq, h, d), dtype=q.dtype, device=dev)
                    P.workspace, P.workspace_bytes = q_rot.data_ptr(), q_rot.numel() * 2
                check(lib.fa_b200_append_kv(C.byref(P), _stream(dev)))
                P.knew = P.vnew = P.rotary_cos = P.rotary_sin = None
                P.workspace, P.workspace_bytes = None, 0
                if q_rot is not None:
                    P.q = q_rot.data_ptr()
                    P.q_batch_stride, P.q_row_stride, P.q_head_stride = q_rot.stride(0), q_rot.stride(1), q_rot.stride(2)
            # else: knew / vnew / rotary stay in the block — the decode kernel appends and rotates itself (one launch)
        else:
            _chk(rotary_cos_ is None, "rotary embedding requires new key/value (reference applies it in the append path)")
        _run(P, OP_KVCACHE, lib.fa_b200_fwd_kvcache, dev)
    return [out, lse]
