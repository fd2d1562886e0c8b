"""Public Python API — signature-compatible with the reference's
vllm_flash_attn/flash_attn_interface.py (v2.5.8.post3): the same seven functions, argument
names, defaults and return conventions (reference lines cited per function).  Forward only:
the reference's extension exports no backward (flash_api.cpp:943-944), so its autograd wrappers
cannot differentiate either; here the functions call the forward op directly.

``flash_attn_cuda`` below is our stand-in for the reference's pybind11 module
``vllm_flash_attn_2_cuda`` (same three ops, same positional arguments).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import vllm_flash_attn_2_cuda as flash_attn_cuda


def _maybe_contiguous(x):
    return x.contiguous() if x is not None and x.stride(-1) != 1 else x


def _flash_attn_forward(q, k, v, dropout_p, softmax_scale, causal, window_size, alibi_slopes,
                        return_softmax, *, out=None):
    """reference flash_attn_interface.py:46-65"""
    q, k, v = [_maybe_contiguous(x) for x in (q, k, v)]
    out, q, k, v, out_padded, softmax_lse, S_dmask, rng_state = flash_attn_cuda.fwd(
        q, k, v, out, alibi_slopes, dropout_p, softmax_scale, causal, window_size[0], window_size[1],
        return_softmax, None)
    return out, q, k, v, out_padded, softmax_lse, S_dmask, rng_state


def _flash_attn_varlen_forward(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k,
                               dropout_p, softmax_scale, causal, window_size, alibi_slopes,
                               return_softmax, block_table, *, out=None):
    """reference flash_attn_interface.py:68-111 (seqused_k=None, zero_tensors=False hard-wired)"""
    q, k, v = [_maybe_contiguous(x) for x in (q, k, v)]
    out, q, k, v, out_padded, softmax_lse, S_dmask, rng_state = flash_attn_cuda.varlen_fwd(
        q, k, v, out, cu_seqlens_q, cu_seqlens_k, None, block_table, alibi_slopes, max_seqlen_q,
        max_seqlen_k, dropout_p, softmax_scale, False, causal, window_size[0], window_size[1],
        return_softmax, None)
    return out, q, k, v, out_padded, softmax_lse, S_dmask, rng_state


def _ret(out, softmax_lse, S_dmask, return_attn_probs):
    return out if not return_attn_probs else (out, softmax_lse, S_dmask)


def flash_attn_qkvpacked_func(qkv, dropout_p=0.0, softmax_scale=None, causal=False,
                              window_size=(-1, -1), alibi_slopes=None, deterministic=False,
                              return_attn_probs=False, *, out=None):
    """qkv: (batch, seqlen, 3, nheads, headdim).  reference :658-713"""
    if softmax_scale is None:
        softmax_scale = qkv.shape[-1] ** (-0.5)
    o, _, _, _, _, lse, S, _ = _flash_attn_forward(
        qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], dropout_p, softmax_scale, causal=causal,
        window_size=window_size, alibi_slopes=alibi_slopes,
        return_softmax=return_attn_probs and dropout_p > 0, out=out)
    return _ret(o, lse, S, return_attn_probs)


def flash_attn_kvpacked_func(q, kv, dropout_p=0.0, softmax_scale=None, causal=False,
                             window_size=(-1, -1), alibi_slopes=None, deterministic=False,
                             return_attn_probs=False, *, out=None):
    """q: (batch, seqlen, nheads, headdim); kv: (batch, seqlen_k, 2, nheads_k, headdim).  reference :716-790"""
    if softmax_scale is None:
        softmax_scale = q.shape[-1] ** (-0.5)
    o, _, _, _, _, lse, S, _ = _flash_attn_forward(
        q, kv[:, :, 0], kv[:, :, 1], dropout_p, softmax_scale, causal=causal, window_size=window_size,
        alibi_slopes=alibi_slopes, return_softmax=return_attn_probs and dropout_p > 0, out=out)
    return _ret(o, lse, S, return_attn_probs)


def flash_attn_func(q, k, v, dropout_p=0.0, softmax_scale=None, causal=False, window_size=(-1, -1),
                    alibi_slopes=None, deterministic=False, return_attn_probs=False, *, out=None):
    """softmax(Q K^T * scale [+ ALiBi] [masked]) V for dense (batch, seqlen, nheads, headdim)
    tensors; MQA/GQA when k/v have fewer heads (query head h uses kv head h // (H/H_k)).
    causal=True aligns the mask to the bottom-right corner; window_size=(left,right) restricts
    query i to keys [i+sk-sq-left, i+sk-sq+right].  Returns out, or (out, softmax_lse, S_dmask)
    if return_attn_probs.  reference :793-867"""
    if softmax_scale is None:
        softmax_scale = q.shape[-1] ** (-0.5)
    o, _, _, _, _, lse, S, _ = _flash_attn_forward(
        q, k, v, dropout_p, softmax_scale, causal=causal, window_size=window_size,
        alibi_slopes=alibi_slopes, return_softmax=return_attn_probs and dropout_p > 0, out=out)
    return _ret(o, lse, S, return_attn_probs)


def flash_attn_varlen_qkvpacked_func(qkv, cu_seqlens, max_seqlen, dropout_p=0.0, softmax_scale=None,
                                     causal=False, window_size=(-1, -1), alibi_slopes=None,
                                     deterministic=False, return_attn_probs=False, *, out=None):
    """qkv: (total, 3, nheads, headdim).  reference :870-932"""
    if softmax_scale is None:
        softmax_scale = qkv.shape[-1] ** (-0.5)
    o, _, _, _, _, lse, S, _ = _flash_attn_varlen_forward(
        qkv[:, 0], qkv[:, 1], qkv[:, 2], cu_seqlens, cu_seqlens, max_seqlen, max_seqlen, dropout_p,
        softmax_scale, causal=causal, window_size=window_size, alibi_slopes=alibi_slopes,
        return_softmax=return_attn_probs and dropout_p > 0, block_table=None, out=out)
    return _ret(o, lse, S, return_attn_probs)


def flash_attn_varlen_kvpacked_func(q, kv, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k,
                                    dropout_p=0.0, softmax_scale=None, causal=False,
                                    window_size=(-1, -1), alibi_slopes=None, deterministic=False,
                                    return_attn_probs=False, *, out=None):
    """q: (total_q, nheads, headdim); kv: (total_k, 2, nheads_k, headdim).  reference :935-1023"""
    if softmax_scale is None:
        softmax_scale = q.shape[-1] ** (-0.5)
    o, _, _, _, _, lse, S, _ = _flash_attn_varlen_forward(
        q, kv[:, 0], kv[:, 1], cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, dropout_p,
        softmax_scale, causal=causal, window_size=window_size, alibi_slopes=alibi_slopes,
        return_softmax=return_attn_probs and dropout_p > 0, block_table=None, out=out)
    return _ret(o, lse, S, return_attn_probs)


def flash_attn_varlen_func(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k,
                           dropout_p=0.0, softmax_scale=None, causal=False, window_size=(-1, -1),
                           alibi_slopes=None, deterministic=False, return_attn_probs=False,
                           block_table=None, *, out=None):
    """Packed variable-length attention: q (total_q, nheads, headdim), k/v (total_k, nheads_k,
    headdim) with int32 cu_seqlens_{q,k} of shape (batch+1,), or — with block_table (batch,
    max_blocks) int32 — k/v as a paged pool (num_blocks, page_block_size, nheads_k, headdim).
    softmax_lse (if returned) is (batch, nheads, max_seqlen_q).  reference :1026-1116"""
    if softmax_scale is None:
        softmax_scale = q.shape[-1] ** (-0.5)
    o, _, _, _, _, lse, S, _ = _flash_attn_varlen_forward(
        q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, dropout_p, softmax_scale,
        causal=causal, window_size=window_size, alibi_slopes=alibi_slopes,
        return_softmax=return_attn_probs and dropout_p > 0, block_table=block_table, out=out)
    return _ret(o, lse, S, return_attn_probs)


def flash_attn_with_kvcache(q, k_cache, v_cache, k=None, v=None, rotary_cos=None, rotary_sin=None,
                            cache_seqlens: Optional[Tuple[int, torch.Tensor]] = None,
                            cache_batch_idx: Optional[torch.Tensor] = None,
                            block_table: Optional[torch.Tensor] = None, softmax_scale=None,
                            causal=False, window_size=(-1, -1), rotary_interleaved=True,
                            alibi_slopes=None, num_splits=0, *, out=None,
                            return_softmax_lse=False):
    """Decode-time attention against a KV cache (reference :1119-1253).

    q (batch, seqlen_q, nheads, headdim); k_cache/v_cache (batch_cache, seqlen_cache, nheads_k,
    headdim), or with block_table (batch, max_blocks) int32 a paged pool (num_blocks,
    page_block_size, nheads_k, headdim) with page_block_size % 16 == 0.  cache_seqlens: int or
    (batch,) int32.  If k/v (batch, seqlen_new, nheads_k, headdim) are given they are appended
    in place at cache_seqlens (rotary applied to k and q when rotary_cos/sin are given) and
    attention runs over the updated cache.  num_splits: 0 = heuristic, 1 = no split.
    Returns out (batch, seqlen_q, nheads, headdim); ``return_softmax_lse`` (keyword-only
    extension found in later vLLM forks, default False) also returns the (batch, nheads, seqlen_q)
    fp32 log-sum-exp, which the reference computes but drops (:1253).
    """
    assert k_cache.stride(-1) == 1, "k_cache must have contiguous last dimension"
    assert v_cache.stride(-1) == 1, "v_cache must have contiguous last dimension"
    q, k, v = [_maybe_contiguous(x) for x in (q, k, v)]
    if softmax_scale is None:
        softmax_scale = q.shape[-1] ** (-0.5)
    if cache_seqlens is not None and isinstance(cache_seqlens, int):
        nb = q.shape[0] if block_table is not None else k_cache.shape[0]
        cache_seqlens = torch.full((nb,), cache_seqlens, dtype=torch.int32, device=k_cache.device)
        cache_seqlens = _maybe_contiguous(cache_seqlens)
    cache_batch_idx = _maybe_contiguous(cache_batch_idx)
    block_table = _maybe_contiguous(block_table)
    out, softmax_lse = flash_attn_cuda.fwd_kvcache(
        q, k_cache, v_cache, k, v, cache_seqlens, rotary_cos, rotary_sin, cache_batch_idx, block_table,
        alibi_slopes, out, softmax_scale, causal, window_size[0], window_size[1], rotary_interleaved,
        num_splits)
    return (out, softmax_lse) if return_softmax_lse else out
