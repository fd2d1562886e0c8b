// fa_append.cu — append new K/V rows into the (paged) KV cache, with rotary embedding on K and Q.
//
// The reference fuses this into the split-KV attention kernel (src/flash_fwd_kernel.h:672-824,
// src/rotary.h:19-145): new rows are written at positions cache_seqlens[b] + i, K is rotated with
// cos/sin taken at its position, Q with cos/sin at cache_seqlens[b] + (causal|local ? qpos : 0).
// Here it is a separate, tiny, purely bandwidth-bound kernel that runs before the attention
// kernel on the same stream (the attention kernel then sees seqlen = cache_seqlens + seqlen_knew),
// which also removes the benign GQA write race the reference documents (:678-680).
// Rotation is done in fp32 and rounded once, like the reference.
#include "fa_common.cuh"
#include "fa_launch.h"
#include "fa_rotary.cuh"

namespace fa {

// grid (rows, heads, b): rows [0, seqlen_knew) are new K/V rows (heads < h_k), rows
// [seqlen_knew, seqlen_knew + seqlen_q) are Q rows to rotate (only when rotary is on).
template <typename T>
__global__ void __launch_bounds__(32) fa_append_kernel(const AppendParams a) {
  const int row = blockIdx.x, head = blockIdx.y, b = blockIdx.z;
  const int tid = threadIdx.x;
  const int cache_len = a.cache_seqlens[b];
  const int half = a.rotary_dim / 2;
  if (row < a.seqlen_knew) {
    if (head >= a.h_k) return;
    const int pos = cache_len + row;
    int64_t koff, voff;
    if (a.block_table) {
      const int pg = min(pos / a.page_block_size, a.max_blocks_per_seq - 1);
      const int blk = a.block_table[(int64_t)b * a.block_table_batch_stride + pg];
      const int in_page = pos - (pos / a.page_block_size) * a.page_block_size;
      koff = (int64_t)blk * a.k_batch_stride + (int64_t)in_page * a.k_row_stride;
      voff = (int64_t)blk * a.v_batch_stride + (int64_t)in_page * a.v_row_stride;
    } else {
      const int kb = a.cache_batch_idx ? a.cache_batch_idx[b] : b;
      if (pos >= a.seqlen_k) return;   // cache full: the reference would write out of bounds
      koff = (int64_t)kb * a.k_batch_stride + (int64_t)pos * a.k_row_stride;
      voff = (int64_t)kb * a.v_batch_stride + (int64_t)pos * a.v_row_stride;
    }
    const T* ksrc = reinterpret_cast<const T*>(a.knew) + (int64_t)b * a.knew_batch_stride + (int64_t)row * a.knew_row_stride + (int64_t)head * a.knew_head_stride;
    const T* vsrc = reinterpret_cast<const T*>(a.vnew) + (int64_t)b * a.vnew_batch_stride + (int64_t)row * a.vnew_row_stride + (int64_t)head * a.vnew_head_stride;
    T* kdst = reinterpret_cast<T*>(a.kcache) + koff + (int64_t)head * a.k_head_stride;
    T* vdst = reinterpret_cast<T*>(a.vcache) + voff + (int64_t)head * a.v_head_stride;
    const int rp = min(pos, max(a.seqlen_ro - 1, 0));
    rotate_row<T>(ksrc, kdst, a.d, a.rotary_dim, a.is_rotary_interleaved,
                  reinterpret_cast<const T*>(a.rotary_cos) + (int64_t)rp * half,
                  reinterpret_cast<const T*>(a.rotary_sin) + (int64_t)rp * half, tid, 32);
    rotate_row<T>(vsrc, vdst, a.d, 0, true, nullptr, nullptr, tid, 32);
  } else {
    const int qpos = row - a.seqlen_knew;
    if (head >= a.h || a.q_rot == nullptr) return;
    const int pos = cache_len + (a.rot_by_qpos ? qpos : 0);
    const int rp = min(pos, max(a.seqlen_ro - 1, 0));
    const T* qsrc = reinterpret_cast<const T*>(a.q) + (int64_t)b * a.q_batch_stride + (int64_t)qpos * a.q_row_stride + (int64_t)head * a.q_head_stride;
    T* qdst = reinterpret_cast<T*>(a.q_rot) + (((int64_t)b * a.seqlen_q + qpos) * a.h + head) * a.d;
    rotate_row<T>(qsrc, qdst, a.d, a.rotary_dim, a.is_rotary_interleaved,
                  reinterpret_cast<const T*>(a.rotary_cos) + (int64_t)rp * half,
                  reinterpret_cast<const T*>(a.rotary_sin) + (int64_t)rp * half, tid, 32);
  }
}

cudaError_t launch_append(const AppendParams& a, bool is_bf16, cudaStream_t stream) {
  const int rows = a.seqlen_knew + (a.q_rot ? a.seqlen_q : 0);
  if (rows == 0) return cudaSuccess;
  dim3 grid(rows, a.q_rot ? max(a.h, a.h_k) : a.h_k, a.b);
  if (is_bf16) fa_append_kernel<__nv_bfloat16><<<grid, 32, 0, stream>>>(a);
  else fa_append_kernel<__half><<<grid, 32, 0, stream>>>(a);
  return cudaGetLastError();
}

// ---- cache writer: token-major K/V rows -> their slots in the paged cache --------------------------------------------
// What a serving engine runs once per step BEFORE attention (vLLM: reshape_and_cache_flash): token t's (h_k, d) row
// goes to cache[slot / page][slot % page], slot = slot_mapping[t]; negative slots (padding tokens) are skipped.
// One warp per (token, kv head), 16-byte chunks; element size 2 (fp16 / bf16).  Not part of the reference repository
// ("the cache-layout / block_table producer", the component on the caller's side of the path).
__global__ void __launch_bounds__(128) fa_reshape_and_cache_kernel(const CacheWriteParams a) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t item = (int64_t)blockIdx.x * 4 + warp;
  if (item >= (int64_t)a.num_tokens * a.h_k) return;
  const int t = (int)(item / a.h_k), hd = (int)(item - (int64_t)t * a.h_k);
  const int64_t slot = a.slot_mapping[t];
  if (slot < 0) return;
  const int64_t blk = slot / a.page, in_page = slot - blk * a.page;
  const uint16_t* ks = reinterpret_cast<const uint16_t*>(a.key) + (int64_t)t * a.key_token_stride + (int64_t)hd * a.key_head_stride;
  const uint16_t* vs = reinterpret_cast<const uint16_t*>(a.value) + (int64_t)t * a.value_token_stride + (int64_t)hd * a.value_head_stride;
  uint16_t* kd = reinterpret_cast<uint16_t*>(a.kcache) + blk * a.k_batch_stride + in_page * a.k_row_stride + (int64_t)hd * a.k_head_stride;
  uint16_t* vd = reinterpret_cast<uint16_t*>(a.vcache) + blk * a.v_batch_stride + in_page * a.v_row_stride + (int64_t)hd * a.v_head_stride;
  for (int c = lane; c < a.d / 8; c += 32) {
    *reinterpret_cast<uint4*>(kd + c * 8) = *reinterpret_cast<const uint4*>(ks + c * 8);
    *reinterpret_cast<uint4*>(vd + c * 8) = *reinterpret_cast<const uint4*>(vs + c * 8);
  }
}

cudaError_t launch_reshape_and_cache(const CacheWriteParams& a, cudaStream_t stream) {
  const int64_t items = (int64_t)a.num_tokens * a.h_k;
  if (items == 0) return cudaSuccess;
  fa_reshape_and_cache_kernel<<<(unsigned)((items + 3) / 4), 128, 0, stream>>>(a);
  return cudaGetLastError();
}

}  // namespace fa
