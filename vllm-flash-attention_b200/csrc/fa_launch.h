// fa_launch.h — host-side launch entry points of the kernels (one per .cu file).
#pragma once
#include "fa_b200.h"
#include <cuda_runtime.h>
#include "fa_common.cuh"

namespace fa {

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) once per (kernel instantiation, device): the
// call costs microseconds, which matters for launch-latency-bound shapes.  Expands inside a
// function template so the flag array is per instantiation.
#define FA_SET_SMEM_ONCE(kern, bytes)                                                         \
  do {                                                                                        \
    static bool fa_attr_done[64] = {false};                                                   \
    int fa_dev = 0;                                                                           \
    cudaGetDevice(&fa_dev);                                                                   \
    if (fa_dev < 0 || fa_dev >= 64 || !fa_attr_done[fa_dev]) {                                \
      cudaError_t fa_e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes); \
      if (fa_e != cudaSuccess) return fa_e;                                                   \
      if (fa_dev >= 0 && fa_dev < 64) fa_attr_done[fa_dev] = true;                            \
    }                                                                                         \
  } while (0)

// fa_decode.cu — rows = seqlen_q*group packed rows must satisfy decode_supported(d, rows).
bool decode_supported(int d, int rows);
cudaError_t launch_decode(const KParams& p, bool is_bf16, cudaStream_t stream);

// fa_tile.cu — feature-complete mma.sync tile kernel, d in {32,64,...,256}.
bool tile_supported(int d);
int tile_block_m(int d);
cudaError_t launch_tile(const KParams& p, bool is_bf16, cudaStream_t stream);

// fa_combine.cu — LSE merge of split partials into (o, lse).
// part_dtype/out_dtype: 0 fp32, 1 bf16, 2 fp16.  Output row (b,h,q) is written to
// o + [cu_q ? cu_q[b]*o_row_stride : b*o_batch_stride] + q*o_row_stride + h*o_head_stride.
struct CombineParams {
  const void* o_part; const float* lse_part; void* o; float* lse;
  int n_splits; int b, h, sq, d;
  int64_t o_batch_stride, o_row_stride, o_head_stride;
  const int* cu_seqlens_q;
  int part_dtype, out_dtype;
  int64_t part_split_stride, lse_split_stride;   // elements between consecutive partials
  int empty_neg_inf;                             // all-empty rows report lse = -inf (partial-result mode)
};
cudaError_t launch_combine(const CombineParams& c, cudaStream_t stream);

// fa_append.cu — append new K/V rows to the cache (+ rotary on K and Q).
struct AppendParams {
  const void* knew; const void* vnew; void* kcache; void* vcache;
  const void* q; void* q_rot;                 // q_rot: (b,seqlen_q,h,d) contiguous, or NULL
  int64_t knew_batch_stride, knew_row_stride, knew_head_stride;
  int64_t vnew_batch_stride, vnew_row_stride, vnew_head_stride;
  int64_t k_batch_stride, k_row_stride, k_head_stride;
  int64_t v_batch_stride, v_row_stride, v_head_stride;
  int64_t q_batch_stride, q_row_stride, q_head_stride;
  int b, h, h_k, d, seqlen_q, seqlen_k, seqlen_knew;
  const int* cache_seqlens; const int* cache_batch_idx;
  const int* block_table; int64_t block_table_batch_stride; int page_block_size, max_blocks_per_seq;
  const void* rotary_cos; const void* rotary_sin; int rotary_dim, seqlen_ro, is_rotary_interleaved;
  int rot_by_qpos;                            // causal or local: q row i rotated at cache_len + i
};
cudaError_t launch_append(const AppendParams& a, bool is_bf16, cudaStream_t stream);
struct CacheWriteParams {
  const void* key; const void* value; void* kcache; void* vcache; const int64_t* slot_mapping;
  int64_t key_token_stride, key_head_stride, value_token_stride, value_head_stride;
  int64_t k_batch_stride, k_row_stride, k_head_stride, v_batch_stride, v_row_stride, v_head_stride;
  int num_tokens, h_k, d, page;
};
cudaError_t launch_reshape_and_cache(const CacheWriteParams& a, cudaStream_t stream);

// fa_prefill_tc5.cu — tcgen05/TMA/TMEM prefill kernel, every head dim 32..256 (multiples of 32).
bool tc5_supported(const KParams& p);
cudaError_t launch_prefill_tc5(const KParams& p, bool is_bf16, int64_t total_q, int64_t total_k, int num_sms,
                               cudaStream_t stream);
cudaError_t launch_exchange_combine(const FaExchangeParams& a, cudaStream_t stream);
size_t tc5_workspace_bytes(const KParams& p, int64_t total_q, int num_sms);

}  // namespace fa
