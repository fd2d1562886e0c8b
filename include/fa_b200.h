/*
 * fa_b200.h — C ABI of the B200-native flash-attention forward library (libfa_b200.so).
 *
 * This is the drop-in boundary: a torch-free shared library whose three entry points
 * replace the three ops the reference exports through pybind11
 * (reference csrc/flash_attn/flash_api.cpp:939-946):
 *
 *   fa_b200_fwd          <-  mha_fwd          (flash_api.cpp:240-420)
 *   fa_b200_varlen_fwd   <-  mha_varlen_fwd   (flash_api.cpp:422-669)
 *   fa_b200_fwd_kvcache  <-  mha_fwd_kvcache  (flash_api.cpp:671-937)
 *
 * All three take one POD parameter block (the analogue of Flash_fwd_params,
 * reference csrc/flash_attn/src/flash.h:51-141): raw device pointers, int64 strides in
 * ELEMENTS, plain ints.  No torch types, no allocation: the caller owns every buffer,
 * including the split-KV workspace whose size fa_b200_workspace_bytes() reports.
 * Kernels are enqueued on the given stream; nothing synchronises the host.
 *
 * Error convention: 0 on success, a negative FA_ERR_* code otherwise; the message
 * (same wording as the reference's TORCH_CHECK text where one exists) is retrievable
 * per thread with fa_b200_last_error().  The Python shim raises RuntimeError from it,
 * matching the reference's c10::Error -> RuntimeError behaviour (flash_api.cpp:16-18).
 */
#ifndef FA_B200_H_
#define FA_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FA_B200_ABI_VERSION 5

enum {
  FA_OK = 0,
  FA_ERR_INVALID = -1,     /* argument validation failed (reference: TORCH_CHECK) */
  FA_ERR_UNSUPPORTED = -2, /* valid in the reference, not built here */
  FA_ERR_CUDA = -3,        /* CUDA runtime / launch error */
  FA_ERR_WORKSPACE = -4    /* workspace too small */
};

/* Which kernel family served the call (written to FaFwdParams.kernel_used). */
enum {
  FA_KERNEL_NONE = 0,
  FA_KERNEL_TC5_PREFILL = 1, /* tcgen05 + TMA + TMEM warp-specialised kernel */
  FA_KERNEL_TILE = 2,        /* mma.sync 64-row tile kernel (all features) */
  FA_KERNEL_DECODE = 3       /* GQA-packed paged split-KV decode kernel */
};

/* Sequence-sharded decode in ONE kernel per rank (fwd_kvcache with FaFwdParams.fused_exchange set): the decode
 * kernel's last CTA per (sequence, kv head) "unit" merges the unit's split-KV partials, pushes the merged fp32 rows
 * into every rank's receive area through peer-mapped pointers (NVLink stores) as FLAGGED words — each fp32 value
 * in one 64-bit word together with the unit's call number, so the data is its own arrival flag: one one-way trip, no
 * system-scope fence, no separate flag — polls the peers' words and writes the final (out, lse): no separate combine or
 * exchange launch, no NCCL call.  No
 * reference analogue (the reference has no multi-GPU path); the algebra is combine_attn_seqk_parallel's
 * (flash_fwd_kernel.h:1151-1221) applied to splits, then to ranks.  All areas live in one symmetric allocation per
 * rank, zeroed once, mapped into every process of the box (e.g. torch symmetric memory):
 *   recv  [2][world][max_units][unit_elems] 64-bit words {fp32 value, uint32 call number}: receive area, double-
 *                                                  buffered by the unit's call parity (8 bytes per value)
 *   flags [world][max_units] uint32                reserved (the flagged words replaced the per-unit flags)
 *   state [max_units] uint32                       this rank's per-unit call numbers
 *   split_counters [max_units] int32               arrival counters of the in-kernel split merge (self-resetting)
 * Every rank must make the same sequence of calls (same b, h_k per call).  Slots are private to a unit, so the
 * batch size may change from call to call. */
#define FA_MAX_PEERS 8
typedef struct FaFusedExchange {
  void* recv[FA_MAX_PEERS];      /* rank r's receive area as mapped into THIS process */
  uint32_t* flags[FA_MAX_PEERS]; /* rank r's flag array as mapped into THIS process */
  uint32_t* state;               /* own */
  int32_t* split_counters;       /* own */
  int32_t rank, world;           /* world <= FA_MAX_PEERS */
  int32_t max_units;             /* capacity: b*h_k of a call must not exceed it */
  int64_t unit_elems;            /* values per slot (8 bytes each): >= seqlen_q*(h/h_k)*(d+1), multiple of 4 */
} FaFusedExchange;

typedef struct FaFwdParams {
  /* ---- tensors (device pointers), strides in elements ---- */
  const void* q;  /* dense (b,sq,h,d) | varlen (total_q,h,d) */
  const void* k;  /* dense (b,sk,hk,d) | varlen (total_k,hk,d) | paged (num_blocks,page,hk,d) */
  const void* v;
  void* o;        /* same shape as q */
  float* softmax_lse; /* (b,h,sq) fp32; varlen: (b,h,max_seqlen_q) (flash_api.cpp:348,578) */
  int64_t q_batch_stride, q_row_stride, q_head_stride;
  int64_t k_batch_stride, k_row_stride, k_head_stride; /* paged: batch stride = per-block stride */
  int64_t v_batch_stride, v_row_stride, v_head_stride;
  int64_t o_batch_stride, o_row_stride, o_head_stride;

  /* ---- sizes ---- */
  int32_t b;         /* batch */
  int32_t seqlen_q;  /* dense: sq; varlen: max_seqlen_q */
  int32_t seqlen_k;  /* dense: sk; varlen: max_seqlen_k; paged kvcache: max_blocks*page */
  int32_t h, h_k, d;
  int32_t is_bf16;   /* 1 = bf16, 0 = fp16 */
  float softmax_scale;

  /* ---- masking ---- */
  int32_t is_causal;
  int32_t window_left, window_right; /* -1 = unbounded */
  const float* alibi_slopes;         /* (h) or (b,h) fp32, or NULL */
  int64_t alibi_slopes_batch_stride; /* 0 when (h) */

  /* ---- varlen / kv-cache ---- */
  const int32_t* cu_seqlens_q; /* (b+1) or NULL */
  const int32_t* cu_seqlens_k; /* (b+1) cumulative, or (b) lengths if !is_seqlens_k_cumulative */
  const int32_t* seqused_k;    /* (b) or NULL */
  int32_t is_seqlens_k_cumulative;
  const int32_t* block_table;  /* (b,max_blocks_per_seq) int32 or NULL */
  int64_t block_table_batch_stride;
  int32_t page_block_size;
  int32_t max_blocks_per_seq;
  int32_t num_blocks;          /* size of the page pool; fwd_kvcache without block_table: batch entries of the cache
                                  (bounds / tensor maps; 0 = unknown, the prefill-sized call then runs on the tile kernel) */
  const int32_t* cache_batch_idx; /* (b) or NULL */
  int64_t total_q, total_k;    /* varlen: rows in q / k (tensor-map bounds) */

  /* ---- append-KV + rotary (fwd_kvcache only) ---- */
  const void* knew; const void* vnew; /* (b,seqlen_knew,hk,d) or NULL */
  int64_t knew_batch_stride, knew_row_stride, knew_head_stride;
  int64_t vnew_batch_stride, vnew_row_stride, vnew_head_stride;
  int32_t seqlen_knew;
  const void* rotary_cos; const void* rotary_sin; /* (seqlen_ro, rotary_dim/2) */
  int32_t rotary_dim, seqlen_ro, is_rotary_interleaved;

  /* ---- split-KV ---- */
  int32_t num_splits;    /* in: 0 = heuristic, >=1 forced; out: splits used (mixed varlen batch: by its decode half) */
  void* workspace;       /* >= fa_b200_workspace_bytes(p) bytes, or NULL when that is 0 */
  size_t workspace_bytes;
  /* fwd_kvcache only.  1 = partial-result mode for a KV range that is one shard of a longer
   * sequence (sequence-sharded decode): `o` is FP32 laid out (b,h,seqlen_q,d) contiguous
   * (strides must say so), softmax_lse is (b,h,seqlen_q) with -inf (not +inf) for rows that see no
   * key in this shard — exactly the per-split partials of the reference's split-KV kernel
   * (flash_api.cpp:212-213, flash_fwd_kernel.h:513-552), ready for fa_b200_combine. */
  int32_t partial_out;
  /* fwd_kvcache, sequence-sharded decode: the cache passed in holds keys [kv_shard_start,
   * kv_shard_start + kv_shard_len) of every sequence and cache_seqlens are the GLOBAL lengths; the kernels
   * use clamp(len - start, 0, kv_shard_len).  kv_shard_len = 0 disables it. */
  int32_t kv_shard_start, kv_shard_len;

  /* ---- tuning / diagnostics ---- */
  int32_t num_sms;       /* 0 = query the device */
  int32_t force_kernel;  /* 0 = auto, else FA_KERNEL_* (tests) */
  int32_t kernel_used;   /* out */
  int32_t launches;      /* out: kernels launched by this call */

  /* in: 1 = the caller guarantees that the first FA_B200_WS_HEAD_BYTES of `workspace` were zero when the buffer was
   * first handed to this library and that nothing but this library has written there since (e.g. it was allocated
   * zeroed and is used only as this library's workspace).  That region holds the arrival counters of the decode
   * kernel's in-kernel split merge, which the kernel leaves at zero; with 0 the library clears them itself with a
   * cudaMemsetAsync on every splitting call. */
  int32_t workspace_clean;

  /* ---- fwd_kvcache, sequence-sharded decode in one kernel (see FaFusedExchange); NULL = off.  Needs kv_shard_len > 0,
   * a decode-shaped call (seqlen_q*(h/h_k) <= 32, 16 above d = 128), no ALiBi, no partial_out; o / softmax_lse
   * receive the FINAL result on every rank. */
  const FaFusedExchange* fused_exchange;
} FaFwdParams;

/* ABI version of the loaded library (== FA_B200_ABI_VERSION). */
int fa_b200_abi_version(void);
/* sizeof(FaFwdParams) as compiled; the binding checks it against its own mirror. */
size_t fa_b200_params_size(void);
/* Thread-local message of the last failing call on this thread ("" if none). */
const char* fa_b200_last_error(void);

/* Split choice for (batch*heads*m_blocks, SMs, n_blocks): the reference's wave-efficiency
 * rule, num_splits_heuristic (flash_api.cpp:158-192). Pure host function. */
int fa_b200_num_splits_heuristic(int batch_nheads_mblocks, int num_sms, int num_n_blocks,
                                 int max_splits);

/* Layout of every workspace: [FA_B200_WS_HEAD_BYTES: split arrival counters, see workspace_clean][payload: split-KV
 * partials or the prefill work list].  16-byte aligned. */
#define FA_B200_WS_HEAD_BYTES 16384
/* Bytes of split-KV workspace the call described by *p needs (0 if it will not split).
 * `op` is 0 = fwd, 1 = varlen_fwd, 2 = fwd_kvcache.  Resolves p->num_splits==0 with the
 * heuristic exactly as the launch will.  Returns (size_t)-1 on invalid arguments. */
size_t fa_b200_workspace_bytes(const FaFwdParams* p, int op);

/* The three ops.  `stream` is a cudaStream_t passed as void*. */
int fa_b200_fwd(FaFwdParams* p, void* stream);
int fa_b200_varlen_fwd(FaFwdParams* p, void* stream);
int fa_b200_fwd_kvcache(FaFwdParams* p, void* stream);

/* Will fa_b200_fwd_kvcache append the call's new K/V rows (knew / vnew, with rotary on K and q when rotary_cos/sin are
 * set) INSIDE its own kernel, as the reference's split kernel does (flash_fwd_kernel.h:672-824)?  1 = yes: pass
 * knew / vnew / rotary_* / seqlen_knew to fa_b200_fwd_kvcache directly (one launch, the rotated q never exists in
 * memory); 0 = no (not a decode shape, sliding window, or kv_shard): run fa_b200_append_kv first; < 0 = invalid. */
int fa_b200_fused_append_supported(const FaFwdParams* p);

/* Append the new K/V rows of a fwd_kvcache call into the (paged) cache at cache_seqlens[b]+i, with
 * rotary embedding on K (and on q, written to p->workspace as (b,seqlen_q,h,d)) when rotary_cos/sin
 * are set — the append half of the reference's fused path (flash_fwd_kernel.h:672-824).  The caller
 * then runs fa_b200_fwd_kvcache with knew/vnew/rotary cleared, seqlen_knew kept, q = rotated q. */
int fa_b200_append_kv(FaFwdParams* p, void* stream);

/* fwd_kvcache for callers that cannot allocate (CUDA-graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  No counterpart in the reference.
 * Same parameters and results as fa_b200_fwd_kvcache, except that p->workspace must be NULL and
 * p->workspace_bytes 0: the split-KV partials (and the prefill work list, when a many-row call lands on
 * the tcgen05 kernel) live in a static device array of FA_B200_STATIC_SCRATCH_BYTES inside the library,
 * which every GPU that loads the library holds until exit.  A call that would need more is rejected with
 * FA_ERR_WORKSPACE and a message stating both sizes; nothing is launched.  The entry point calls no
 * cudaMalloc / cudaFree / allocating routine.  All calls share the array (it keeps the last call's
 * contents; nothing in it has to start at zero): calls that overlap on the GPU race on it, so
 * serialise them on one stream or use fa_b200_fwd_kvcache with separate workspaces. */
#define FA_B200_STATIC_SCRATCH_BYTES (4u << 20)
int fa_b200_fwd_kvcache_static_scratch(FaFwdParams* p, void* stream);

/* Cache writer of the caller's side (no counterpart in the reference repository; vLLM: reshape_and_cache_flash, SURVEY 8f
 * rank 4): row t of key / value, (num_tokens, h_k, d) fp16 or bf16 with the given element strides, is copied to
 * cache[slot / page][slot % page] with slot = slot_mapping[t] (int64, device); negative slots are skipped.  The caches are
 * (num_blocks, page_block_size, h_k, d) with the given strides — the layout the three ops read. */
int fa_b200_reshape_and_cache(const void* key, const void* value, void* kcache, void* vcache, const int64_t* slot_mapping,
                              int num_tokens, int h_k, int d, int page_block_size, int64_t num_blocks,
                              int64_t key_token_stride, int64_t key_head_stride, int64_t value_token_stride, int64_t value_head_stride,
                              int64_t k_batch_stride, int64_t k_row_stride, int64_t k_head_stride,
                              int64_t v_batch_stride, int64_t v_row_stride, int64_t v_head_stride, void* stream);

/* Merge `n` partial results into o (rows,d) + lse (rows): lse = logsumexp_s lse_s,
 * o = sum_s exp(lse_s-lse) o_s — the algebra of combine_attn_seqk_parallel
 * (flash_fwd_kernel.h:1092-1250); used for the in-GPU split-KV merge and for the cross-GPU
 * (out,lse) merge after an allgather.  Partial s lives at o_part + s*part_stride (rows x d
 * elements, row-major) and lse_part + s*lse_stride (rows floats); strides are in ELEMENTS, so a
 * packed per-rank record [o (rows*d) | lse (rows)] of fp32 can be merged in place with
 * part_stride = lse_stride = rows*(d+1).  part_dtype / out_dtype: 0 = fp32, 1 = bf16, 2 = fp16.
 * Rows whose partials are all empty (-inf) get o = 0 and lse = +inf (flash_fwd_kernel.h:1156,1164). */
int fa_b200_combine(const void* o_part, const float* lse_part, void* o, float* lse,
                    int32_t n, int64_t rows, int32_t d, int64_t part_stride, int64_t lse_stride,
                    int32_t part_dtype, int32_t out_dtype, void* stream);

/* Cross-GPU merge of sequence-sharded decode partials in one kernel per rank (no reference analogue:
 * the reference has no multi-GPU path; the algebra is combine_attn_seqk_parallel's,
 * flash_fwd_kernel.h:1151-1221, lifted from splits to ranks).  Every rank pushes its packed fp32 record
 * [o (b,h,seqlen_q,d) | lse (b,h,seqlen_q)] into all ranks' receive areas through peer-mapped pointers
 * (NVLink stores), publishes a sequence number, waits for the peers' and LSE-combines.  recv[r] / flags[r]
 * are rank r's receive area ([2][world][record_elems] fp32) and flag array ([world][FA_XCHG_CTAS] uint32,
 * zeroed once) as mapped into THIS process (e.g. torch symmetric memory buffer_ptrs); state is this rank's own
 * per-CTA call counters ([FA_XCHG_CTAS] uint32, zeroed once).  All ranks must make the same sequence of calls. */
#define FA_XCHG_CTAS 64
typedef struct FaExchangeParams {
  void* recv[FA_MAX_PEERS];
  uint32_t* flags[FA_MAX_PEERS];
  const void* local_record;  /* this rank's record, 16-byte aligned */
  uint32_t* state;
  int32_t rank, world;
  int32_t b, h, seqlen_q, d; /* d % 4 == 0 */
  int64_t record_elems;      /* fp32 elements per slot: >= b*h*seqlen_q*(d+1), multiple of 4 */
  void* o;                   /* (b,seqlen_q,h,d) by the strides below */
  float* softmax_lse;        /* (b,h,seqlen_q) or NULL */
  int64_t o_batch_stride, o_row_stride, o_head_stride;
  int32_t out_dtype;         /* 0 = fp32, 1 = bf16, 2 = fp16 */
} FaExchangeParams;
int fa_b200_exchange_combine(const FaExchangeParams* p, void* stream);
size_t fa_b200_exchange_params_size(void);
size_t fa_b200_fused_exchange_size(void);   /* sizeof(FaFusedExchange) as compiled */

#ifdef __cplusplus
}
#endif
#endif /* FA_B200_H_ */
