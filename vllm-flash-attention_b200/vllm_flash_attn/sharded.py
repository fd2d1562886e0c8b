"""Multi-GPU partitioning of the attention forward path (one process per GPU, torch.distributed).

The reference has no multi-GPU attention path ("replicas only"); these helpers add
the two shardings BASELINE.json's north_star names, both built on the single-GPU C ABI:

* **head sharding** (tensor-parallel style, no collective): KV heads — each with its H/H_k query
  heads — are independent, so rank r simply owns a contiguous slice of them (`head_shard`).
* **KV-sequence sharding for long-context decode**: rank r owns keys [r*L, (r+1)*L) of every
  sequence in its own page pool, computes a partial `(out_r fp32, lse_r)` with the query replicated,
  the ranks exchange the packed records with ONE collective (NCCL all-gather over NVLink/NVSwitch)
  and every rank merges them with the LSE-combine kernel:
      lse = logsumexp_r lse_r,   out = sum_r exp(lse_r - lse) * out_r
  — the algebra of combine_attn_seqk_parallel (csrc/flash_attn/src/flash_fwd_kernel.h:1151-1221)
  lifted one level (`flash_attn_with_kvcache_seq_sharded`).

Nothing here touches the CPU oracle; the compute steps are libfa_b200.so kernels.  The two
compute hooks (`_local_fn`, `_merge_fn`) exist so that the host-side plumbing (range arithmetic,
record layout, collective) can be exercised by world_size-2 gloo tests on a CPU-only box.
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple

import torch

__all__ = ["head_shard", "seq_shard_range", "seq_shard_lengths", "partial_record_numel",
           "kvcache_partial", "merge_partials", "flash_attn_with_kvcache_seq_sharded"]


# --------------------------------------------------------------------------------------------
# range arithmetic (pure host logic)
def head_shard(rank: int, world: int, nheads: int, nheads_k: int) -> Tuple[slice, slice]:
    """(query-head slice, kv-head slice) owned by `rank`.  KV heads are split contiguously and each
    keeps its g = H/H_k query heads (query head h uses kv head h // g, flash_fwd_kernel.h:126), so a
    rank's q[..., qs, :], k[..., ks, :], v[..., ks, :] form a self-contained attention problem."""
    if nheads % nheads_k != 0:
        raise ValueError("Number of heads in key/value must divide number of heads in query")
    if nheads_k % world != 0:
        raise ValueError(f"nheads_k={nheads_k} is not divisible by world size {world}")
    g = nheads // nheads_k
    per = nheads_k // world
    return slice(rank * per * g, (rank + 1) * per * g), slice(rank * per, (rank + 1) * per)


def seq_shard_range(rank: int, world: int, max_seqlen_k: int, page_block_size: int = 16) -> Tuple[int, int]:
    """Key range [start, end) of every sequence that `rank` owns: equal page-aligned chunks of
    L = ceil(max_seqlen_k / world) rounded up to a whole page."""
    per = -(-max_seqlen_k // world)
    per = -(-per // page_block_size) * page_block_size
    return rank * per, (rank + 1) * per


def seq_shard_lengths(cache_seqlens: torch.Tensor, start: int, end: int) -> torch.Tensor:
    """Per-sequence number of keys inside [start, end): clamp(len - start, 0, end - start), int32."""
    return (cache_seqlens.to(torch.int64) - start).clamp_(0, end - start).to(torch.int32)


def partial_record_numel(b: int, h: int, sq: int, d: int) -> int:
    """fp32 elements of one rank's packed record [o (b,h,sq,d) | lse (b,h,sq)], padded to 16 B."""
    n = b * h * sq * (d + 1)
    return (n + 3) // 4 * 4


# --------------------------------------------------------------------------------------------
# compute steps (CUDA kernels through the C ABI)
def kvcache_partial(q, k_cache, v_cache, cache_seqlens, block_table=None, *, softmax_scale=None,
                    causal=False, window_size=(-1, -1), alibi_slopes=None, num_splits=0,
                    buf: Optional[torch.Tensor] = None):
    """Attention of q against ONE SHARD of the KV sequence: returns (buf, out32, lse) where `buf` is
    the packed fp32 record (see partial_record_numel), out32 its (b,h,sq,d) view and lse its (b,h,sq)
    view, with lse = -inf for rows that see no key in this shard."""
    from . import vllm_flash_attn_2_cuda as ext
    b, sq, h, d = q.shape
    if buf is None:
        buf = torch.empty(partial_record_numel(b, h, sq, d), dtype=torch.float32, device=q.device)
    rec = b * h * sq * (d + 1)
    if softmax_scale is None:
        softmax_scale = d ** -0.5
    out32, lse = ext.fwd_kvcache(q, k_cache, v_cache, None, None, cache_seqlens, None, None, None, block_table,
                                 alibi_slopes, None, softmax_scale, causal, window_size[0], window_size[1],
                                 True, num_splits, _partial_buf=buf[:rec])
    return buf, out32, lse


def merge_partials(records: torch.Tensor, b: int, h: int, sq: int, d: int, out_dtype: torch.dtype):
    """Merge n packed records (n, partial_record_numel) into out (b,sq,h,d) `out_dtype` and
    lse (b,h,sq) fp32 with the LSE-combine kernel (fa_b200_combine)."""
    from ._cabi import check, lib
    n, rec = records.shape
    rows = b * h * sq
    assert records.dtype == torch.float32 and records.is_contiguous() and rec >= rows * (d + 1)
    out = torch.empty((b, h, sq, d), dtype=out_dtype, device=records.device)
    lse = torch.empty((b, h, sq), dtype=torch.float32, device=records.device)
    code = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2}[out_dtype]
    with torch.cuda.device(records.device):
        check(lib.fa_b200_combine(records.data_ptr(), records.data_ptr() + rows * d * 4, out.data_ptr(), lse.data_ptr(),
                                  n, rows, d, rec, rec, 0, code, torch.cuda.current_stream(records.device).cuda_stream))
    return out.transpose(1, 2), lse


# --------------------------------------------------------------------------------------------
def flash_attn_with_kvcache_seq_sharded(q, k_cache, v_cache, cache_seqlens, block_table=None, *,
                                        shard_start: int, shard_end: int, group=None, softmax_scale=None,
                                        causal=False, alibi_slopes=None, num_splits=0, return_softmax_lse=False,
                                        _local_fn: Optional[Callable] = None, _merge_fn: Optional[Callable] = None):
    """Decode attention with the KV sequence sharded over the ranks of `group`.

    q (b, seqlen_q, h, d) is replicated.  This rank's k_cache/v_cache (+ block_table) hold keys
    [shard_start, shard_end) of every sequence (local row 0 = global key shard_start; see
    seq_shard_range); `cache_seqlens` are the GLOBAL lengths.  Every rank returns the full result.

    seqlen_q must be 1 (or causal=False): with the last query token every key of the sequence is
    visible, so no cross-rank mask is needed.  ALiBi is rejected because its bias
    depends on the global key position.
    """
    import torch.distributed as dist
    b, sq, h, d = q.shape
    if sq != 1 and causal:
        raise RuntimeError("sequence-sharded attention supports seqlen_q == 1 or causal=False")
    if alibi_slopes is not None:
        raise RuntimeError("sequence-sharded attention does not support ALiBi")
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    local_lens = seq_shard_lengths(cache_seqlens, shard_start, shard_end)
    rec = partial_record_numel(b, h, sq, d)
    gathered = torch.empty((world, rec), dtype=torch.float32, device=q.device)
    rank = dist.get_rank(group) if world > 1 else 0
    mine = gathered[rank]                       # the local kernel writes straight into its slot
    local = _local_fn or (lambda buf: kvcache_partial(q, k_cache, v_cache, local_lens, block_table,
                                                       softmax_scale=softmax_scale, causal=False,
                                                       num_splits=num_splits, buf=buf))
    local(mine)
    if world > 1:
        # ONE collective: every rank's (out fp32 | lse) record, in place (slot r of `gathered`)
        dist.all_gather_into_tensor(gathered.view(-1), mine, group=group)
    merge = _merge_fn or merge_partials
    out, lse = merge(gathered, b, h, sq, d, q.dtype)
    return (out, lse) if return_softmax_lse else out
