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
           "kvcache_partial", "merge_partials", "flash_attn_with_kvcache_seq_sharded", "PeerExchange"]


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
                    buf: Optional[torch.Tensor] = None, shard: Optional[Tuple[int, int]] = None):
    """Attention of q against ONE SHARD of the KV sequence: returns (buf, out32, lse) where `buf` is
    the packed fp32 record (see partial_record_numel), out32 its (b,h,sq,d) view and lse its (b,h,sq)
    view, with lse = -inf for rows that see no key in this shard.  With shard=(start, end) the
    cache holds keys [start, end) of every sequence and `cache_seqlens` are the GLOBAL lengths (the
    kernels clamp them: no extra launches)."""
    from . import vllm_flash_attn_2_cuda as ext
    b, sq, h, d = q.shape
    if buf is None:
        buf = torch.empty(partial_record_numel(b, h, sq, d), dtype=torch.float32, device=q.device)
    rec = b * h * sq * (d + 1)
    if softmax_scale is None:
        softmax_scale = d ** -0.5
    out32, lse = ext.fwd_kvcache(q, k_cache, v_cache, None, None, cache_seqlens, None, None, None, block_table,
                                 alibi_slopes, None, softmax_scale, causal, window_size[0], window_size[1],
                                 True, num_splits, _partial_buf=buf[:rec], _kv_shard=shard)
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
_XCHG_CTAS = 64            # FA_XCHG_CTAS (include/fa_b200.h)


class PeerExchange:
    """Peer-memory state for the fused cross-GPU merge (fa_b200_exchange_combine): one symmetric
    allocation per rank — receive area [2][world][record], flag array [world][64], state [64], and the local
    staging record the decode kernel writes — mapped into every process of the box by torch symmetric
    memory (NVLink P2P).  Create it once (collectively, all ranks of `group`) for the largest
    (b, h, seqlen_q, d) that will be merged; `merge()` then costs one kernel launch per decode step and
    is CUDA-graph replayable.  All ranks must call merge() the same number of times."""

    @classmethod
    def for_decode(cls, max_batch: int, h: int, h_k: int, d: int, device, group=None, seqlen_q: int = 1):
        """Exchange sized for decode steps of up to `max_batch` sequences with h / h_k heads of dimension d: both
        the two-kernel record and the areas of the ONE-kernel path (fused_exchange)."""
        return cls(partial_record_numel(max_batch, h, seqlen_q, d), device, group,
                   fused_units=max_batch * h_k, fused_unit_elems=(seqlen_q * (h // h_k) * (d + 1) + 3) // 4 * 4)

    def __init__(self, max_record_elems: int, device, group=None, *, fused_units: int = 0, fused_unit_elems: int = 0):
        import torch.distributed as dist
        single = not dist.is_initialized() or dist.get_world_size(group) == 1
        self.group = None if single else (group if group is not None else dist.group.WORLD)
        self.world = 1 if single else dist.get_world_size(self.group)
        self.rank = 0 if single else dist.get_rank(self.group)
        if self.world > 8:
            raise ValueError("PeerExchange supports at most 8 ranks (one NVSwitch box)")
        self.record = (max_record_elems + 3) // 4 * 4
        recv_elems = 2 * self.world * self.record
        self.off_flags = recv_elems                       # fp32/uint32 element offsets inside the allocation
        self.off_state = self.off_flags + 8 * _XCHG_CTAS  # flags [FA_MAX_PEERS][FA_XCHG_CTAS]
        self.off_local = self.off_state + _XCHG_CTAS      # state [FA_XCHG_CTAS]
        # areas of the one-kernel path (include/fa_b200.h, FaFusedExchange); absent when fused_units == 0
        self.fused_units, self.fused_unit_elems = int(fused_units), (int(fused_unit_elems) + 3) // 4 * 4
        self.off_frecv = (self.off_local + self.record + 3) // 4 * 4
        # receive area of the one-kernel path: flagged lines, 8 bytes per fp32 value (value + the call's sequence number)
        self.off_fflags = self.off_frecv + 2 * (2 * self.world * self.fused_units * self.fused_unit_elems)
        self.off_fstate = self.off_fflags + 8 * self.fused_units
        self.off_fcount = self.off_fstate + self.fused_units
        total = self.off_fcount + self.fused_units
        self._fx = None
        if single:                                        # one rank: an ordinary allocation, same protocol
            self.buf = torch.zeros(total, dtype=torch.float32, device=device)
            self.handle = None
            self.ptrs = [self.buf.data_ptr()]
        else:
            import torch.distributed._symmetric_memory as symm_mem
            self.buf = symm_mem.empty(total, dtype=torch.float32, device=device)
            self.buf.zero_()
            self.handle = symm_mem.rendezvous(self.buf, self.group)
            torch.cuda.synchronize(device)
            self.handle.barrier()                         # every rank's flags/state are zero before any push
            self.ptrs = [int(p) for p in self.handle.buffer_ptrs]
        self.local_record = self.buf[self.off_local:self.off_local + self.record]

    def fused_params(self, b: int, h: int, h_k: int, sq: int, d: int):
        """FaFusedExchange for a call of this shape, or None when the one-kernel areas are absent or too small."""
        if self.fused_units == 0 or b * h_k > self.fused_units or sq * (h // h_k) * (d + 1) > self.fused_unit_elems:
            return None
        if self._fx is None:
            from ._cabi import FaFusedExchange
            fx = FaFusedExchange()
            for r in range(self.world):
                fx.recv[r] = self.ptrs[r] + 4 * self.off_frecv
                fx.flags[r] = self.ptrs[r] + 4 * self.off_fflags
            fx.state = self.ptrs[self.rank] + 4 * self.off_fstate
            fx.split_counters = self.ptrs[self.rank] + 4 * self.off_fcount
            fx.rank, fx.world = self.rank, self.world
            fx.max_units, fx.unit_elems = self.fused_units, self.fused_unit_elems
            self._fx = fx
        return self._fx

    def merge(self, b: int, h: int, sq: int, d: int, out_dtype: torch.dtype, rec_elems: Optional[int] = None):
        """LSE-merge the records all ranks have staged in `local_record` (layout of
        partial_record_numel).  Returns (out (b,sq,h,d) out_dtype, lse (b,h,sq) fp32)."""
        from ._cabi import FaExchangeParams, check, lib
        import ctypes as C
        rec = rec_elems or partial_record_numel(b, h, sq, d)
        if rec > self.record:
            raise ValueError(f"record of {rec} fp32 exceeds the PeerExchange capacity {self.record}")
        dev = self.buf.device
        out = torch.empty((b, sq, h, d), dtype=out_dtype, device=dev)
        lse = torch.empty((b, h, sq), dtype=torch.float32, device=dev)
        P = FaExchangeParams()
        for r in range(self.world):
            P.recv[r] = self.ptrs[r]
            P.flags[r] = self.ptrs[r] + 4 * self.off_flags
        P.local_record = self.ptrs[self.rank] + 4 * self.off_local
        P.state = self.ptrs[self.rank] + 4 * self.off_state
        P.rank, P.world = self.rank, self.world
        P.b, P.h, P.seqlen_q, P.d = b, h, sq, d
        P.record_elems = self.record
        P.o, P.softmax_lse = out.data_ptr(), lse.data_ptr()
        P.o_batch_stride, P.o_row_stride, P.o_head_stride = out.stride(0), out.stride(1), out.stride(2)
        P.out_dtype = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2}[out_dtype]
        check(lib.fa_b200_exchange_combine(C.byref(P), torch.cuda.current_stream(dev).cuda_stream))
        return out, lse


# --------------------------------------------------------------------------------------------
def flash_attn_with_kvcache_seq_sharded(q, k_cache, v_cache, cache_seqlens, block_table=None, *,
                                        shard_start: int, shard_end: int, group=None, softmax_scale=None,
                                        causal=False, alibi_slopes=None, num_splits=0, return_softmax_lse=False,
                                        exchange: Optional["PeerExchange"] = None, one_kernel: Optional[bool] = None,
                                        _local_fn: Optional[Callable] = None, _merge_fn: Optional[Callable] = None):
    """Decode attention with the KV sequence sharded over the ranks of `group`.

    q (b, seqlen_q, h, d) is replicated.  This rank's k_cache/v_cache (+ block_table) hold keys
    [shard_start, shard_end) of every sequence (local row 0 = global key shard_start; see
    seq_shard_range); `cache_seqlens` are the GLOBAL lengths.  Every rank returns the full result.

    With `exchange` (a PeerExchange created once for the group) the records travel by peer-memory
    stores instead of an NCCL all-gather.  If the exchange was created with the one-kernel areas
    (PeerExchange.for_decode) and they fit the call, the whole step is ONE kernel launch per rank: the decode
    kernel merges its own splits, pushes the merged rows to the peers, waits for theirs and writes the final
    result (`one_kernel=False` forces the older form: decode + split combine + a push / wait / combine kernel).

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
    shard = (shard_start, shard_end)
    rec = partial_record_numel(b, h, sq, d)
    if exchange is not None and _local_fn is None and one_kernel is not False:
        fx = exchange.fused_params(b, h, k_cache.shape[2], sq, d)
        if fx is not None:
            from . import vllm_flash_attn_2_cuda as ext
            scale = d ** -0.5 if softmax_scale is None else softmax_scale
            out, lse = ext.fwd_kvcache(q, k_cache, v_cache, None, None, cache_seqlens, None, None, None, block_table, None, None,
                                       scale, False, -1, -1, True, num_splits, _kv_shard=shard, _fused_exchange=fx)
            return (out, lse) if return_softmax_lse else out
        if one_kernel:
            raise ValueError("one_kernel=True, but the PeerExchange has no one-kernel areas large enough for this call")
    if exchange is not None:
        if _local_fn is not None:
            _local_fn(exchange.local_record[:rec])
        else:
            kvcache_partial(q, k_cache, v_cache, cache_seqlens, block_table, softmax_scale=softmax_scale, causal=False,
                            num_splits=num_splits, buf=exchange.local_record[:rec], shard=shard)
        out, lse = exchange.merge(b, h, sq, d, q.dtype, rec)
        return (out, lse) if return_softmax_lse else out
    gathered = torch.empty((world, rec), dtype=torch.float32, device=q.device)
    rank = dist.get_rank(group) if world > 1 else 0
    mine = gathered[rank]                       # the local kernel writes straight into its slot
    local = _local_fn or (lambda buf: kvcache_partial(q, k_cache, v_cache, cache_seqlens, block_table,
                                                       softmax_scale=softmax_scale, causal=False,
                                                       num_splits=num_splits, buf=buf, shard=shard))
    local(mine)
    if world > 1:
        # ONE collective: every rank's (out fp32 | lse) record, in place (slot r of `gathered`)
        dist.all_gather_into_tensor(gathered.view(-1), mine, group=group)
    merge = _merge_fn or merge_partials
    out, lse = merge(gathered, b, h, sq, d, q.dtype)
    return (out, lse) if return_softmax_lse else out
