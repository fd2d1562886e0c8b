"""TEST INFRASTRUCTURE — CPU oracle for the flash-attention forward path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product (vllm_flash_attn -> libfa_b200.so) never does.

It restates, in plain PyTorch on CPU, the reference's own ground truth for this path:

* ``attention_ref``            <- /root/reference/tests/test_flash_attn.py:208-288
* ``construct_local_mask``     <- tests/test_flash_attn.py:178-205
* ``attn_bias_from_alibi_slopes`` <- tests/test_flash_attn.py:29-51
* ``generate_random_padding_mask`` <- tests/test_flash_attn.py:54-67
* ``generate_block_kvcache``   <- tests/test_flash_attn.py:2079-2103 (``_generate_block_kvcache``)
* ``flash_blockwise_ref``      <- the kernel's arithmetic (csrc/flash_attn/src/softmax.h:66-186,
  flash_fwd_kernel.h:281-402): blockwise online softmax in fp32, exp2 with folded scale,
  P rounded to the I/O dtype before P.V, row-sum taken on the unrounded P, LSE = max*scale+ln(sum).
* ``combine_ref``              <- flash_fwd_kernel.h:1151-1221 (split-KV / cross-GPU LSE merge)

Parity pin: tests/golden/*.pt were produced by tests/golden/make_golden.py, which executes the
reference's *own* ``attention_ref`` (loaded from /root/reference in the build container) on
seeded inputs; tests/test_oracle.py checks this restatement against those vectors bit for bit.
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch


def construct_local_mask(seqlen_q, seqlen_k, window_size=(-1, -1), query_padding_mask=None,
                         key_padding_mask=None, device=None):
    """True = masked out.  Bottom-right aligned (tests/test_flash_attn.py:178-205)."""
    row = torch.arange(seqlen_q, device=device, dtype=torch.long).view(-1, 1)
    col = torch.arange(seqlen_k, device=device, dtype=torch.long)
    sk = seqlen_k if key_padding_mask is None else key_padding_mask.sum(-1).view(-1, 1, 1, 1)
    sq = seqlen_q if query_padding_mask is None else query_padding_mask.sum(-1).view(-1, 1, 1, 1)
    if window_size[0] < 0:
        return col > row + sk - sq + window_size[1]
    sk_t = torch.full_like(col, seqlen_k) if key_padding_mask is None else sk
    return torch.logical_or(col > torch.minimum(row + sk - sq + window_size[1], sk_t),
                            col < row + sk - sq - window_size[0])


def attn_bias_from_alibi_slopes(slopes, seqlen_q, seqlen_k, query_padding_mask=None,
                                key_padding_mask=None, causal=False):
    """ALiBi bias, (b,h,1|sq,sk) fp32 (tests/test_flash_attn.py:29-51)."""
    batch, nheads = slopes.shape
    device = slopes.device
    slopes = slopes.view(batch, nheads, 1, 1)
    if causal:
        return torch.arange(-seqlen_k + 1, 1, device=device, dtype=torch.float32) * slopes
    row = torch.arange(seqlen_q, device=device, dtype=torch.long).view(-1, 1)
    col = torch.arange(seqlen_k, device=device, dtype=torch.long)
    sk = seqlen_k if key_padding_mask is None else key_padding_mask.sum(-1).view(-1, 1, 1, 1)
    sq = seqlen_q if query_padding_mask is None else query_padding_mask.sum(-1).view(-1, 1, 1, 1)
    rel = torch.abs(row + sk - sq - col)
    return -slopes * rel.to(dtype=slopes.dtype)


def generate_random_padding_mask(max_seqlen, batch_size, device, mode="random"):
    """tests/test_flash_attn.py:54-67."""
    assert mode in ("full", "random", "third")
    if mode == "full":
        lengths = torch.full((batch_size, 1), max_seqlen, device=device, dtype=torch.int32)
    elif mode == "random":
        lengths = torch.randint(max(1, max_seqlen - 20), max_seqlen + 1, (batch_size, 1), device=device)
    else:
        lengths = torch.randint(max_seqlen // 3, max_seqlen + 1, (batch_size, 1), device=device)
    return torch.arange(max_seqlen, device=device).expand(batch_size, max_seqlen) < lengths


def attention_ref(q, k, v, query_padding_mask=None, key_padding_mask=None, attn_bias=None,
                  causal=False, window_size=(-1, -1), upcast=True, reorder_ops=False):
    """Plain softmax(QK^T*scale + bias + mask) V  (tests/test_flash_attn.py:208-288).

    q (b,sq,h,d); k,v (b,sk,hk,d).  Returns (out (b,sq,h,d) in q.dtype, attention (b,h,sq,sk)).
    upcast=False/reorder_ops=True is the reference's "same math in low precision" error yardstick.
    Dropout arguments of the reference are omitted: the reference build compiles dropout out
    (setup.py:204, flash_api.cpp:115-117).
    """
    if causal:
        window_size = (window_size[0], 0)
    dtype_og = q.dtype
    if upcast:
        q, k, v = q.float(), k.float(), v.float()
    seqlen_q, seqlen_k = q.shape[1], k.shape[1]
    g = q.shape[2] // k.shape[2]
    k = k.repeat_interleave(g, dim=2)
    v = v.repeat_interleave(g, dim=2)
    d = q.shape[-1]
    if not reorder_ops:
        scores = torch.einsum("bthd,bshd->bhts", q / math.sqrt(d), k)
    else:
        scores = torch.einsum("bthd,bshd->bhts", q, k / math.sqrt(d))
    if key_padding_mask is not None:
        scores.masked_fill_(~key_padding_mask.view(key_padding_mask.shape[0], 1, 1, -1), float("-inf"))
    local_mask = None
    if window_size[0] >= 0 or window_size[1] >= 0:
        local_mask = construct_local_mask(seqlen_q, seqlen_k, window_size, query_padding_mask,
                                          key_padding_mask, q.device)
        scores.masked_fill_(local_mask, float("-inf"))
    if attn_bias is not None:
        scores = scores + attn_bias
    attention = torch.softmax(scores, dim=-1).to(v.dtype)
    if local_mask is not None:
        attention = attention.masked_fill(torch.all(local_mask, dim=-1, keepdim=True), 0.0)
    if query_padding_mask is not None:
        attention = attention.masked_fill(~query_padding_mask.view(query_padding_mask.shape[0], 1, -1, 1), 0.0)
    output = torch.einsum("bhts,bshd->bthd", attention, v)
    if query_padding_mask is not None:
        output.masked_fill_(~query_padding_mask.view(query_padding_mask.shape[0], -1, 1, 1), 0.0)
    return output.to(dtype=dtype_og), attention.to(dtype=dtype_og)


def lse_ref(q, k, key_padding_mask=None, query_padding_mask=None, attn_bias=None, causal=False,
            window_size=(-1, -1), softmax_scale=None):
    """fp32 log-sum-exp of the scaled, masked scores, (b,h,sq).  Fully masked rows give +inf,
    the kernel's convention (softmax.h:178-179)."""
    if causal:
        window_size = (window_size[0], 0)
    q, k = q.float(), k.float()
    g = q.shape[2] // k.shape[2]
    k = k.repeat_interleave(g, dim=2)
    scale = softmax_scale if softmax_scale is not None else q.shape[-1] ** -0.5
    scores = torch.einsum("bthd,bshd->bhts", q, k) * scale
    if key_padding_mask is not None:
        scores.masked_fill_(~key_padding_mask.view(key_padding_mask.shape[0], 1, 1, -1), float("-inf"))
    if window_size[0] >= 0 or window_size[1] >= 0:
        m = construct_local_mask(q.shape[1], k.shape[1], window_size, query_padding_mask,
                                 key_padding_mask, q.device)
        scores.masked_fill_(m, float("-inf"))
    if attn_bias is not None:
        scores = scores + attn_bias
    lse = torch.logsumexp(scores, dim=-1)
    return torch.where(torch.isinf(lse) & (lse < 0), torch.full_like(lse, float("inf")), lse)


def generate_block_kvcache(seqlen_k, paged_kv_block_size, batch_size, nheads_k, d, device, dtype,
                           generator=None):
    """Random paged cache + permuted block table (tests/test_flash_attn.py:2079-2103).
    Returns k_cache (b,S,hk,d), v_cache, block_table (b,nb) int32, k_cache_paged, v_cache_paged,
    num_blocks; S = ceil(seqlen_k/page)*page; the pool is 3x over-provisioned."""
    num_blocks = math.ceil(seqlen_k / paged_kv_block_size) * batch_size * 3
    k_paged = torch.randn(num_blocks, paged_kv_block_size, nheads_k, d, device=device, dtype=dtype,
                          generator=generator)
    v_paged = torch.randn(num_blocks, paged_kv_block_size, nheads_k, d, device=device, dtype=dtype,
                          generator=generator)
    perm = torch.randperm(num_blocks, dtype=torch.int32, device=device, generator=generator)
    block_table = perm.view(batch_size, num_blocks // batch_size)
    k_cache = gather_paged(k_paged, block_table)[:, :seqlen_k]
    v_cache = gather_paged(v_paged, block_table)[:, :seqlen_k]
    return k_cache, v_cache, block_table, k_paged, v_paged, num_blocks


def gather_paged(cache_paged, block_table):
    """(num_blocks,page,hk,d) + (b,nb) -> contiguous (b, nb*page, hk, d).  Row r of sequence b
    lives in block block_table[b][r // page] at slot r % page (csrc/flash_attn/src/utils.h:297-315)."""
    b, nb = block_table.shape
    g = cache_paged[block_table.to(torch.long).flatten()]
    return g.view(b, nb * cache_paged.shape[1], cache_paged.shape[2], cache_paged.shape[3])


def visible_range(i, seqlen_q, seqlen_k, causal, window_size):
    """[lo, hi) of keys visible to query row i (mask.h:171-172)."""
    wl, wr = window_size
    if causal:
        wr = 0
    shift = i + seqlen_k - seqlen_q
    lo = 0 if wl < 0 else max(0, shift - wl)
    hi = seqlen_k if wr < 0 else min(seqlen_k, shift + 1 + wr)
    return lo, max(hi, lo) if hi >= 0 else lo


def flash_blockwise_ref(q, k, v, softmax_scale=None, causal=False, window_size=(-1, -1),
                        alibi_slopes=None, block_n=128, split_range=None, round_p=True):
    """Kernel-arithmetic model for ONE sequence: q (sq,h,d), k,v (sk,hk,d) -> out fp32 (sq,h,d)
    *unrounded*, lse (h,sq) fp32.  Walks K/V in blocks of block_n with the reference's online
    softmax (softmax.h:135-186): scores kept unscaled, p = exp2(s*scale*log2e - max*scale*log2e),
    row-sum in fp32 on unrounded p, P rounded to q.dtype before P.V (flash_fwd_kernel.h:323,382),
    O rescaled by exp2((m_old-m_new)*scale*log2e), final O/sum, LSE = max*scale + ln(sum).
    ALiBi: bias = -slope*|i+sk-sq-j| added to the scaled score (mask.h:181-183 equivalent).
    split_range=(k0,k1) restricts to keys [k0,k1) -> a split-KV partial (empty => O=0, lse=-inf,
    flash_fwd_kernel.h:513-552).  Small sizes only (python loop over blocks)."""
    sq, h, d = q.shape
    sk, hk, _ = k.shape
    g = h // hk
    scale = softmax_scale if softmax_scale is not None else d ** -0.5
    scale_log2 = scale * math.log2(math.e)
    dt = q.dtype
    qf = q.float()
    kf = k.float().repeat_interleave(g, dim=1)
    vf = v.float().repeat_interleave(g, dim=1)
    wl, wr = window_size
    if causal:
        wr = 0
    rows = torch.arange(sq).view(1, sq, 1)
    k0, k1 = (0, sk) if split_range is None else split_range
    m = torch.full((h, sq), float("-inf"))
    l = torch.zeros(h, sq)
    o = torch.zeros(h, sq, d)
    for n0 in range(k0, k1, block_n):
        n1 = min(n0 + block_n, k1)
        s = torch.einsum("thd,shd->hts", qf, kf[n0:n1])          # unscaled
        cols = torch.arange(n0, n1).view(1, 1, -1)
        shift = rows + sk - sq
        if alibi_slopes is not None:
            s = s + (-(alibi_slopes.float().view(h, 1, 1) / scale) * (shift - cols).abs())
        masked = torch.zeros(1, sq, n1 - n0, dtype=torch.bool)
        if wr >= 0:
            masked = masked | (cols > shift + wr)
        if wl >= 0:
            masked = masked | (cols < shift - wl)
        s = s.masked_fill(masked, float("-inf"))
        m_new = torch.maximum(m, s.amax(dim=-1))
        m_safe = torch.where(torch.isinf(m_new), torch.zeros_like(m_new), m_new)
        p = torch.exp2(s * scale_log2 - (m_safe * scale_log2).unsqueeze(-1))
        alpha = torch.exp2((torch.where(torch.isinf(m), m_safe, m) - m_safe) * scale_log2)
        alpha = torch.where(torch.isinf(m), torch.zeros_like(alpha), alpha)
        l = l * alpha + p.sum(dim=-1)
        pr = p.to(dt).float() if round_p else p
        o = o * alpha.unsqueeze(-1) + torch.einsum("hts,shd->htd", pr, vf[n0:n1])
        m = m_new
    empty = (l == 0) | torch.isnan(l)
    inv = torch.where(empty, torch.ones_like(l), 1.0 / l)
    out = (o * inv.unsqueeze(-1)).permute(1, 0, 2).contiguous()
    lse_empty = float("-inf") if split_range is not None else float("inf")
    lse = torch.where(empty, torch.full_like(l, lse_empty),
                      torch.where(torch.isinf(m), torch.zeros_like(m), m) * scale + torch.log(l))
    return out, lse


def combine_ref(o_part, lse_part):
    """Merge split partials: o_part (S,rows,d), lse_part (S,rows) fp32 -> (o (rows,d) fp32,
    lse (rows)).  lse = logsumexp_s(lse_s) (+inf if every split is -inf), o = sum_s
    exp(lse_s - lse) o_s  (flash_fwd_kernel.h:1151-1221)."""
    o_part = o_part.float()
    lse_max = lse_part.amax(dim=0)
    lse_max_safe = torch.where(torch.isinf(lse_max) & (lse_max < 0), torch.zeros_like(lse_max), lse_max)
    ssum = torch.exp(lse_part - lse_max_safe).sum(dim=0)
    lse = torch.where((ssum == 0) | torch.isnan(ssum), torch.full_like(ssum, float("inf")),
                      torch.log(ssum) + lse_max_safe)
    w = torch.exp(lse_part - lse)            # exp(-inf) = 0 for the all-empty case
    o = (w.unsqueeze(-1) * o_part).sum(dim=0)
    return o, lse


def attention_ref_varlen(q, k, v, cu_seqlens_q, cu_seqlens_k, causal=False, window_size=(-1, -1),
                         alibi_slopes=None, upcast=True, reorder_ops=False):
    """Per-sequence attention_ref over packed (total,h,d) tensors; returns packed out and an
    (b,h,max_sq) fp32 lse (padded with 0 beyond each sequence)."""
    b = len(cu_seqlens_q) - 1
    out = torch.empty_like(q)
    max_sq = int(max(cu_seqlens_q[i + 1] - cu_seqlens_q[i] for i in range(b)))
    lse = torch.zeros(b, q.shape[1], max_sq)
    for i in range(b):
        q0, q1 = int(cu_seqlens_q[i]), int(cu_seqlens_q[i + 1])
        k0, k1 = int(cu_seqlens_k[i]), int(cu_seqlens_k[i + 1])
        if q1 == q0:
            continue
        bias = None
        if alibi_slopes is not None:
            sl = alibi_slopes if alibi_slopes.dim() == 2 else alibi_slopes.unsqueeze(0)
            sl = sl[i:i + 1] if sl.shape[0] > 1 else sl
            # sliding-window causal is 'local' to the kernels (flash_api.cpp:119-126): ALiBi in its |i - j| form (mask.h:183)
            bias = attn_bias_from_alibi_slopes(sl.float(), q1 - q0, k1 - k0, causal=causal and window_size[0] < 0)
        qi, ki, vi = q[q0:q1].unsqueeze(0), k[k0:k1].unsqueeze(0), v[k0:k1].unsqueeze(0)
        if k1 == k0:
            out[q0:q1] = 0
            lse[i, :, :q1 - q0] = float("inf")
            continue
        o, _ = attention_ref(qi, ki, vi, attn_bias=bias, causal=causal, window_size=window_size,
                             upcast=upcast, reorder_ops=reorder_ops)
        out[q0:q1] = o[0]
        l_i = lse_ref(qi, ki, attn_bias=bias, causal=causal, window_size=window_size)[0]
        if bias is not None and causal and window_size[0] < 0:   # sliding-window causal is 'local' to the kernel: |i-j| form, no shift
            # kernel convention: causal ALiBi bias is slope*j (csrc/flash_attn/src/mask.h:181), the
            # test helper's is slope*(j-(sk-1)) (tests/test_flash_attn.py:38-39): same out, LSE
            # differs by the row-constant slope*(sk-1)
            l_i = l_i + sl.float().view(-1, 1).to(l_i.device) * (k1 - k0 - 1)
        lse[i, :, :q1 - q0] = l_i
    return out, lse
