// fa_combine.cu — log-sum-exp merge of split-KV partials.
//
// Same algebra as combine_attn_seqk_parallel (reference src/flash_fwd_kernel.h:1092-1250):
//   lse = logsumexp_s(lse_s)   (all splits empty (-inf) => lse = +inf, O = 0; :1156,:1164)
//   O   = sum_s exp(lse_s - lse) * O_s
// Also used as the cross-GPU merge of per-rank (out, lse) after an allgather.
//
// Design: one warp per output row; lanes first own splits (max / sum by shuffles), then own
// 4-float column groups and walk the splits in order with float4 loads — 512 B coalesced per
// split for d=128, fixed summation order (deterministic).
#include "fa_common.cuh"
#include "fa_launch.h"

namespace fa {

template <typename PT> __device__ __forceinline__ float4 load4(const PT* p);
template <> __device__ __forceinline__ float4 load4<float>(const float* p) {
  return *reinterpret_cast<const float4*>(p);
}
template <> __device__ __forceinline__ float4 load4<__nv_bfloat16>(const __nv_bfloat16* p) {
  uint2 u = *reinterpret_cast<const uint2*>(p);
  __nv_bfloat162 a = *reinterpret_cast<__nv_bfloat162*>(&u.x), b = *reinterpret_cast<__nv_bfloat162*>(&u.y);
  return make_float4(__low2float(a), __high2float(a), __low2float(b), __high2float(b));
}
template <> __device__ __forceinline__ float4 load4<__half>(const __half* p) {
  uint2 u = *reinterpret_cast<const uint2*>(p);
  __half2 a = *reinterpret_cast<__half2*>(&u.x), b = *reinterpret_cast<__half2*>(&u.y);
  return make_float4(__low2float(a), __high2float(a), __low2float(b), __high2float(b));
}

template <typename PT, int OUT_DT>
__global__ void __launch_bounds__(128) fa_combine_kernel(const CombineParams c) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t rows = (int64_t)c.b * c.h * c.sq;
  const int64_t row = (int64_t)blockIdx.x * 4 + warp;
  if (row >= rows) return;
  const int q = (int)(row % c.sq);
  const int hh = (int)((row / c.sq) % c.h);
  const int bb = (int)(row / ((int64_t)c.sq * c.h));
  int64_t obase;
  if (c.cu_seqlens_q) {
    int s0 = c.cu_seqlens_q[bb];
    if (q >= c.cu_seqlens_q[bb + 1] - s0) return;   // padded row of a varlen batch
    obase = (int64_t)s0 * c.o_row_stride;
  } else {
    obase = (int64_t)bb * c.o_batch_stride;
  }
  obase += (int64_t)q * c.o_row_stride + (int64_t)hh * c.o_head_stride;

  const int S = c.n_splits;   // <= 128: each lane owns up to 4 splits
  float ls[4];
  float mx = -INFINITY;
#pragma unroll
  for (int i = 0; i < 4; i++) {
    int s = lane + 32 * i;
    ls[i] = s < S ? c.lse_part[(int64_t)s * c.lse_split_stride + row] : -INFINITY;
    mx = fmaxf(mx, ls[i]);
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
  const float mx_safe = mx == -INFINITY ? 0.f : mx;
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < 4; i++) sum += expf(ls[i] - mx_safe);
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
  const bool empty = !(sum > 0.f);
  const float lse = empty ? (c.empty_neg_inf ? -INFINITY : INFINITY) : logf(sum) + mx_safe;
  if (lane == 0 && c.lse) c.lse[row] = lse;
  float wgt[4];
#pragma unroll
  for (int i = 0; i < 4; i++) wgt[i] = empty ? 0.f : expf(ls[i] - lse);

  const PT* part = reinterpret_cast<const PT*>(c.o_part);
  for (int col0 = 0; col0 < c.d; col0 += 128) {   // warp-uniform trip count: the shuffles need all lanes
    const int col = col0 + lane * 4;
    const bool live = col < c.d;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int i = 0; i < 4; i++) {
      for (int sl = 0; sl < 32 && i * 32 + sl < S; sl++) {
        const int s = i * 32 + sl;
        float w = __shfl_sync(0xffffffffu, wgt[i], sl);
        if (live) {
          float4 x = load4<PT>(part + (int64_t)s * c.part_split_stride + row * c.d + col);
          acc.x += w * x.x; acc.y += w * x.y; acc.z += w * x.z; acc.w += w * x.w;
        }
      }
    }
    if (!live) continue;
    if constexpr (OUT_DT == 0) {
      *reinterpret_cast<float4*>(reinterpret_cast<float*>(c.o) + obase + col) = acc;
    } else if constexpr (OUT_DT == 1) {
      uint2 pk; pk.x = Pack2<__nv_bfloat16>::pack(acc.x, acc.y); pk.y = Pack2<__nv_bfloat16>::pack(acc.z, acc.w);
      *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(c.o) + obase + col) = pk;
    } else {
      uint2 pk; pk.x = Pack2<__half>::pack(acc.x, acc.y); pk.y = Pack2<__half>::pack(acc.z, acc.w);
      *reinterpret_cast<uint2*>(reinterpret_cast<__half*>(c.o) + obase + col) = pk;
    }
  }
}

template <typename PT>
static cudaError_t launch_combine_p(const CombineParams& c, cudaStream_t stream) {
  const int64_t rows = (int64_t)c.b * c.h * c.sq;
  if (rows == 0) return cudaSuccess;
  dim3 grid((unsigned)((rows + 3) / 4));
  switch (c.out_dtype) {
    case 0: fa_combine_kernel<PT, 0><<<grid, 128, 0, stream>>>(c); break;
    case 1: fa_combine_kernel<PT, 1><<<grid, 128, 0, stream>>>(c); break;
    case 2: fa_combine_kernel<PT, 2><<<grid, 128, 0, stream>>>(c); break;
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

cudaError_t launch_combine(const CombineParams& c, cudaStream_t stream) {
  if (c.n_splits < 1 || c.n_splits > 128 || c.d % 4 != 0) return cudaErrorInvalidValue;
  switch (c.part_dtype) {
    case 0: return launch_combine_p<float>(c, stream);
    case 1: return launch_combine_p<__nv_bfloat16>(c, stream);
    case 2: return launch_combine_p<__half>(c, stream);
  }
  return cudaErrorInvalidValue;
}

}  // namespace fa
