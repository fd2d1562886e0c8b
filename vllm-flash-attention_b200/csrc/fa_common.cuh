// fa_common.cuh — device-side parameter block and sm_100a PTX wrappers shared by the kernels.
//
// KParams is the resolved form of FaFwdParams (include/fa_b200.h) after the host glue
// (fa_api.cu) has applied the reference's normalisation rules (flash_api.cpp:289-294,
// :119-126).  It plays the role of Flash_fwd_params (reference src/flash.h:51-141).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace fa {

struct KParams {
  const void* q; const void* k; const void* v; void* o;
  float* lse;             // (b,h,seqlen_q) fp32
  float* o_accum;         // (splits,b,h,seqlen_q,d) fp32 — split partials
  float* lse_accum;       // (splits,b,h,seqlen_q)
  int64_t q_batch_stride, q_row_stride, q_head_stride;
  int64_t k_batch_stride, k_row_stride, k_head_stride;
  int64_t v_batch_stride, v_row_stride, v_head_stride;
  int64_t o_batch_stride, o_row_stride, o_head_stride;
  int b, seqlen_q, seqlen_k, h, h_k, d, group;  // group = h / h_k
  float scale, scale_log2;                      // softmax_scale, softmax_scale*log2(e)
  int is_causal, is_local;                      // local: window applies and not pure causal
  int window_left, window_right;                // normalised as set_params_fprop does
  const float* alibi_slopes; int64_t alibi_batch_stride;
  const int* cu_seqlens_q; const int* cu_seqlens_k; const int* seqused_k;
  int is_seqlens_k_cumulative;
  const int* block_table; int64_t block_table_batch_stride; int page_block_size;
  int max_blocks_per_seq; int num_blocks;
  const int* cache_batch_idx;
  int seqlen_knew;                              // keys appended this call (already in cache)
  int num_splits;
  int lse_seq_stride;                           // seqlen_q used for lse/accum row indexing
  int part_seq_stride;                          // query positions per (split, b, h) in the decode kernel's split partials (= lse_seq_stride,
                                                // except for the decode half of a mixed batch, whose sequences have only a few positions)
  int min_rows, max_rows;                        // serve only sequences with min_rows <= seqlen_q*group <= max_rows (mixed batches)
  int kv_shard_start, kv_shard_len;               // sequence-sharded decode: this rank holds keys [start, start+len) of every sequence (len 0 = off)
  int2* work_items; int* work_header;           // tcgen05 prefill: cost-sorted (sequence, m_block) list + {draw counter, count}; NULL = static order
  long long* trace;                             // FA_TC5_TRACE builds only: per-event clock64 stamps of one CTA
  int single_page;                              // tcgen05 prefill over a NON-paged cache: each cache entry is one page of seqlen_k rows,
                                                // block_table (if any) is cache_batch_idx with one column, else block = sequence index
  int partial;                                  // write fp32 partials (o_accum/lse_accum, -inf for empty rows) even with 1 split
  // ---- decode kernel: in-kernel split merge and cross-GPU exchange (fa_decode.cu, "fused tail") ----
  int* split_counters;                          // [b*h_k] arrival counters, zero before the launch, reset by the kernel; NULL: the
                                                // split partials are merged by a separate fa_combine_kernel launch
  float* xchg_recv[8]; unsigned* xchg_flags[8]; // sequence-sharded decode in ONE kernel: peer-mapped receive areas / flag arrays
  unsigned* xchg_state;                         // this rank's per-unit call numbers
  int xchg_rank, xchg_world, xchg_max_units;    // xchg_world == 0: no exchange
  int64_t xchg_unit_elems;                      // fp32 per (parity, source rank, unit) slot
  // ---- decode kernel: in-kernel append of the call's new K/V rows (+ rotary on K and q), fa_decode.cu ----
  const void* knew; const void* vnew;           // (b, seqlen_knew, h_k, d) or NULL: rows are already in the cache
  int64_t knew_batch_stride, knew_row_stride, knew_head_stride;
  int64_t vnew_batch_stride, vnew_row_stride, vnew_head_stride;
  const void* rotary_cos; const void* rotary_sin;   // (seqlen_ro, rotary_dim / 2) or NULL
  int rotary_dim, seqlen_ro, rotary_interleaved, rot_by_qpos;
  // ---- decode kernel: balanced ranges over all units (fa_decode_streamk_kernel) ----
  int sk_grid;                                  // > 0: launch that kernel with this many CTAs
  int sk_min_per_cta;                           // chunks per CTA at least
  int sk_rows;                                  // rows per piece slot of the partial pool (sk_grid + b*h_k slots)
  int wide_cta;                                 // decode, d = 128, instead of 4 warps x 3 stages: 1 = 6 x 2 (one wave of long pieces), 2 = 4 x 2, three CTAs per SM (many waves)
};
constexpr int FA_KP_MAX_PEERS = 8;

// Per-sequence geometry — the analogue of BlockInfo (reference src/block_info.h:11-42).
struct SeqInfo {
  int sq, sk;            // actual lengths
  int64_t q_off, k_off, v_off;  // element offsets of row 0 (paged: k_off/v_off unused)
  int kv_batch;          // batch index into a non-paged cache
  __device__ __forceinline__ SeqInfo(const KParams& p, int bidb) {
    if (p.cu_seqlens_q) {
      int s0 = p.cu_seqlens_q[bidb];
      sq = p.cu_seqlens_q[bidb + 1] - s0;
      q_off = (int64_t)s0 * p.q_row_stride;
    } else {
      sq = p.seqlen_q;
      q_off = (int64_t)bidb * p.q_batch_stride;
    }
    kv_batch = p.cache_batch_idx ? p.cache_batch_idx[bidb] : bidb;
    if (p.cu_seqlens_k) {
      if (p.is_seqlens_k_cumulative) {
        int s0 = p.cu_seqlens_k[bidb];
        sk = p.cu_seqlens_k[bidb + 1] - s0;
        k_off = (int64_t)s0 * p.k_row_stride; v_off = (int64_t)s0 * p.v_row_stride;
      } else {
        sk = p.cu_seqlens_k[bidb] + p.seqlen_knew;
        k_off = (int64_t)kv_batch * p.k_batch_stride; v_off = (int64_t)kv_batch * p.v_batch_stride;
      }
    } else {
      sk = p.seqlen_k;
      k_off = (int64_t)kv_batch * p.k_batch_stride; v_off = (int64_t)kv_batch * p.v_batch_stride;
    }
    if (p.seqused_k) sk = p.seqused_k[bidb];
    if (p.kv_shard_len > 0) sk = min(max(sk - p.kv_shard_start, 0), p.kv_shard_len);   // global length -> keys held locally
  }
};

// Output row offsets (o uses the same row indexing as q).
__device__ __forceinline__ int64_t o_seq_offset(const KParams& p, int bidb) {
  return p.cu_seqlens_q ? (int64_t)p.cu_seqlens_q[bidb] * p.o_row_stride
                        : (int64_t)bidb * p.o_batch_stride;
}

// ---------------------------------------------------------------------------------------
// dtype helpers
template <typename T> struct Pack2;
template <> struct Pack2<__nv_bfloat16> {
  static __device__ __forceinline__ uint32_t pack(float lo, float hi) {
    uint32_t r; asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo)); return r;
  }
};
template <> struct Pack2<__half> {
  static __device__ __forceinline__ uint32_t pack(float lo, float hi) {
    uint32_t r; asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo)); return r;
  }
};

__device__ __forceinline__ float fast_exp2(float x) {
  float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

// system-scope flag traffic of the cross-GPU exchanges (peer-mapped memory over NVLink)
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// ---- flagged 16-byte lines (the "LL" form NCCL uses for small messages): two fp32 values, each packed with the call's
// sequence number into one 64-bit word.  A naturally aligned 64-bit store is single-copy atomic, so the receiver that
// sees the sequence number in a word also sees its value: the data is its own arrival flag, and a cross-GPU transfer
// costs ONE one-way trip instead of data + system-scope release fence (a round trip) + flag (another one-way trip).
__device__ __forceinline__ void st_flagged(float* line, float a, float b, uint32_t n) {
  const unsigned long long x = (unsigned long long)__float_as_uint(a) | ((unsigned long long)n << 32);
  const unsigned long long y = (unsigned long long)__float_as_uint(b) | ((unsigned long long)n << 32);
  asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(line), "l"(x), "l"(y) : "memory");
}
#ifndef FA_XCHG_TIMEOUT_NS
#define FA_XCHG_TIMEOUT_NS 120000000000ll   /* a peer may be late by a module load, a graph capture or a host stall; a MISSING peer still ends in a launch failure, not a hung GPU */
#endif
// wait until *flag has reached call number n (wrap-safe); bounded
__device__ __forceinline__ void wait_flag_sys(const unsigned* flag, unsigned n) {
  long long t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  while ((int)(ld_acquire_sys(flag) - n) < 0) {
    long long t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    if (t1 - t0 > FA_XCHG_TIMEOUT_NS) __trap();
  }
}
// Poll one flagged line until both words carry sequence number n (bounded like wait_flag_sys).
__device__ __forceinline__ float2 ld_flagged(const float* line, uint32_t n) {
  unsigned long long x, y;
  long long t0 = 0;
  for (uint32_t spins = 0;; spins++) {
    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(x), "=l"(y) : "l"(line) : "memory");
    if ((uint32_t)(x >> 32) == n && (uint32_t)(y >> 32) == n) break;
    if ((spins & 1023u) == 1023u) {
      long long now; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
      if (t0 == 0) t0 = now;
      else if (now - t0 > FA_XCHG_TIMEOUT_NS) __trap();
    }
  }
  return make_float2(__uint_as_float((uint32_t)x), __uint_as_float((uint32_t)y));
}

// ---------------------------------------------------------------------------------------
// cp.async (LDGSTS) 16-byte, zero-fill when !pred
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool pred) {
  int sz = pred ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N));
}

// ldmatrix / mma.sync (legacy tensor path — used by the HBM-bound decode kernel and the
// feature-complete tile kernel; the compute-bound prefill kernel uses tcgen05 below)
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
template <typename T> __device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1);
template <> __device__ __forceinline__ void mma16816<__nv_bfloat16>(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <> __device__ __forceinline__ void mma16816<__half>(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

}  // namespace fa
