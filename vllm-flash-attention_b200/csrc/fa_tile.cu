// fa_tile.cu — feature-complete tile kernel on mma.sync (64 query rows per CTA, 4 warps).
//
// Covers every combination the reference's two kernels accept (compute_attn_1rowblock,
// src/flash_fwd_kernel.h:27-470, and compute_attn_1rowblock_splitkv, :474-1053): dense / varlen /
// paged KV / cache_batch_idx / seqused_k, causal / sliding window / ALiBi, MHA/GQA/MQA, split-KV,
// head dims 32..256 (multiples of 32, flash_api.cpp:135-137), fp16 / bf16.  It is the safety net
// under the two B200-specialised kernels: fa_prefill_tc5.cu (tcgen05, all head dims) takes the
// compute-bound prefill shapes and fa_decode.cu the HBM-bound decode shapes.
//
// One CTA = 64 query rows of one (sequence, head[, split]); each warp owns 16 rows and walks the
// visible KV range in 64-key tiles that all four warps share through a double-buffered cp.async
// pipeline.  Paged K/V rows are resolved per 16 B segment (block_table lookups hit L1).
#include "fa_common.cuh"
#include "fa_warp_attn.cuh"
#include "fa_launch.h"

namespace fa {

template <int D> struct TileCfg {
  // d >= 192: 64-key tiles would leave room for one CTA per SM only (160 KB at d=256); 32-key tiles keep two
  // (measured: d=256 causal 184 -> 232 TFLOP/s, d=192 160 -> 205)
  static constexpr int BM = 64, BN = D >= 192 ? 32 : 64, NW = 4;
  static constexpr int Q_BYTES = BM * D * 2;
  static constexpr int KV_BYTES = BN * D * 2;
  static constexpr int SMEM = Q_BYTES + 4 * KV_BYTES;   // K,V x 2 stages
};

template <typename T, int D>
__global__ void __launch_bounds__(128) fa_tile_kernel(const __grid_constant__ KParams p) {
  using Cfg = TileCfg<D>;
  constexpr int BM = Cfg::BM, BN = Cfg::BN, CPR = D / 8;
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t q_s = smem_u32(smem);
  const uint32_t kv_s = q_s + Cfg::Q_BYTES;

  const int m_blocks = (p.seqlen_q + BM - 1) / BM;
  const int m_block = blockIdx.x % m_blocks, split = blockIdx.x / m_blocks;
  const int head = blockIdx.y, bidb = blockIdx.z;
  const int kvh = head / p.group;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const SeqInfo si(p, bidb);
  const int sq = si.sq, sk = si.sk;
  const int m0 = m_block * BM;
  if (m0 >= sq || sq * p.group < p.min_rows || sq * p.group > p.max_rows) return;
  const int m_last = min(m0 + BM, sq) - 1;

  // visible KV tiles for this row block, then this split's share
  const RowWindow w_first = row_window(p, m0, sq, sk), w_last = row_window(p, m_last, sq, sk);
  const int t_lo = max(w_first.lo, 0) / BN;
  const int t_hi = (max(w_last.hi, 0) + BN - 1) / BN;
  const int n_tiles = max(t_hi - t_lo, 0);
  const int per_split = (n_tiles + p.num_splits - 1) / p.num_splits;
  const int t_begin = t_lo + split * per_split;
  const int t_end = min(t_hi, t_begin + per_split);

  // ---- Q tile ----
  {
    const T* qb = reinterpret_cast<const T*>(p.q) + si.q_off + (int64_t)head * p.q_head_stride;
    for (int idx = threadIdx.x; idx < BM * CPR; idx += 128) {
      int r = idx / CPR, c = idx % CPR;
      bool ok = m0 + r < sq;
      cp_async16(q_s + swz<D>(r, c), qb + (int64_t)(ok ? m0 + r : m0) * p.q_row_stride + c * 8, ok);
    }
    cp_async_commit();
  }

  const T* kbase = reinterpret_cast<const T*>(p.k) + (int64_t)kvh * p.k_head_stride;
  const T* vbase = reinterpret_cast<const T*>(p.v) + (int64_t)kvh * p.v_head_stride;
  const bool paged = p.block_table != nullptr;
  const int* table = paged ? p.block_table + (int64_t)bidb * p.block_table_batch_stride : nullptr;
  if (!paged) { kbase += si.k_off; vbase += si.v_off; }
  const int page = p.page_block_size;

  auto issue = [&](int t, int buf) {
    const int row0 = t * BN;
    const uint32_t ks = kv_s + buf * 2 * Cfg::KV_BYTES, vs = ks + Cfg::KV_BYTES;
    for (int idx = threadIdx.x; idx < BN * CPR; idx += 128) {
      int r = idx / CPR, c = idx % CPR;
      int row = row0 + r;
      bool ok = row < sk;
      int rr = ok ? row : 0;
      int64_t koff, voff;
      if (paged) {
        int pg = min(rr / page, p.max_blocks_per_seq - 1);
        int blk = table[pg], in_page = rr - (rr / page) * page;
        koff = (int64_t)blk * p.k_batch_stride + (int64_t)in_page * p.k_row_stride;
        voff = (int64_t)blk * p.v_batch_stride + (int64_t)in_page * p.v_row_stride;
      } else {
        koff = (int64_t)rr * p.k_row_stride; voff = (int64_t)rr * p.v_row_stride;
      }
      cp_async16(ks + swz<D>(r, c), kbase + koff + c * 8, ok);
      cp_async16(vs + swz<D>(r, c), vbase + voff + c * 8, ok);
    }
  };

  // ---- per-thread row metadata ----
  int row_lo[2], row_hi[2], row_shift[2];
  const bool has_alibi = p.alibi_slopes != nullptr;
  const float slope2 = has_alibi ? p.alibi_slopes[bidb * p.alibi_batch_stride + head] * 1.4426950408889634f : 0.f;
#pragma unroll
  for (int r = 0; r < 2; r++) {
    int qpos = min(m0 + warp * 16 + (lane >> 2) + r * 8, sq - 1);
    RowWindow w = row_window(p, qpos, sq, sk);
    row_lo[r] = w.lo; row_hi[r] = w.hi; row_shift[r] = qpos + sk - sq;
  }
  const int hi_min = w_first.hi, lo_max = w_last.lo;

  WarpAttn<T, D, 1, BN> attn;
  attn.init();

  if (t_begin < t_end) issue(t_begin, 0);
  cp_async_commit();
  const uint32_t my_q = q_s + warp * 16 * D * 2;
  for (int t = t_begin; t < t_end; t++) {
    const int buf = (t - t_begin) & 1;
    if (t + 1 < t_end) issue(t + 1, buf ^ 1);
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();
    const uint32_t ks = kv_s + buf * 2 * Cfg::KV_BYTES;
    const int col0 = t * BN;
    const bool need_mask = has_alibi || (col0 + BN > hi_min) || (col0 < lo_max);
    attn.step(my_q, ks, ks + Cfg::KV_BYTES, p.scale_log2, need_mask,
              [&](int row, int col, float x) -> float {
                const int r2 = (row >> 3) & 1;
                const int j = col0 + col;
                if (has_alibi) x += p.is_causal ? slope2 * (float)j                // src/mask.h:181
                                                : -slope2 * fabsf((float)(row_shift[r2] - j));
                return (j >= row_lo[r2] && j < row_hi[r2]) ? x : -INFINITY;
              });
    __syncthreads();
  }
  cp_async_wait<0>();
  __syncthreads();   // Q tile landed even when there were no KV tiles; smem reusable
  attn.finish_sums();

  // ---- epilogue ----
  float inv[2], lse[2];
#pragma unroll
  for (int r = 0; r < 2; r++) {
    const bool empty = !(attn.l[0][r] > 0.f);
    inv[r] = empty ? 0.f : 1.f / attn.l[0][r];
    lse[r] = empty ? ((p.num_splits > 1 || p.partial) ? -INFINITY : INFINITY)
                   : attn.m[0][r] * 0.6931471805599453f + __logf(attn.l[0][r]);
  }
  const int row_base = m0 + warp * 16;
  if ((p.num_splits > 1 || p.partial)) {
#pragma unroll
    for (int r = 0; r < 2; r++) {
      int qpos = row_base + (lane >> 2) + r * 8;
      if (qpos < sq) {
        int64_t orow = (((int64_t)split * p.b + bidb) * p.h + head) * p.lse_seq_stride + qpos;
        float* dst = p.o_accum + orow * D + (lane & 3) * 2;
#pragma unroll
        for (int j = 0; j < D / 8; j++)
          *reinterpret_cast<float2*>(dst + j * 8) = make_float2(attn.o[0][j][2 * r] * inv[r], attn.o[0][j][2 * r + 1] * inv[r]);
        if ((lane & 3) == 0) p.lse_accum[orow] = lse[r];
      }
    }
  } else {
    // stage the warp's 16 x D tile through its slice of the (now free) Q region, then 16 B stores
    uint8_t* stage = smem + warp * 16 * D * 2;
#pragma unroll
    for (int r = 0; r < 2; r++) {
      int lr = (lane >> 2) + r * 8;
#pragma unroll
      for (int j = 0; j < D / 8; j++)
        *reinterpret_cast<uint32_t*>(stage + swz<D>(lr, j) + (lane & 3) * 4) =
            Pack2<T>::pack(attn.o[0][j][2 * r] * inv[r], attn.o[0][j][2 * r + 1] * inv[r]);
    }
    __syncwarp();
    T* ob = reinterpret_cast<T*>(p.o) + o_seq_offset(p, bidb) + (int64_t)head * p.o_head_stride;
    for (int idx = lane; idx < 16 * CPR; idx += 32) {
      int lr = idx / CPR, c = idx % CPR;
      int qpos = row_base + lr;
      if (qpos < sq)
        *reinterpret_cast<uint4*>(ob + (int64_t)qpos * p.o_row_stride + c * 8) =
            *reinterpret_cast<const uint4*>(stage + swz<D>(lr, c));
    }
    if (p.lse && (lane & 3) == 0) {
#pragma unroll
      for (int r = 0; r < 2; r++) {
        int qpos = row_base + (lane >> 2) + r * 8;
        if (qpos < sq) p.lse[((int64_t)bidb * p.h + head) * p.lse_seq_stride + qpos] = lse[r];
      }
    }
  }
}

template <typename T, int D>
static cudaError_t launch_tile_t(const KParams& p, cudaStream_t stream) {
  using Cfg = TileCfg<D>;
  auto kern = fa_tile_kernel<T, D>;
  FA_SET_SMEM_ONCE(kern, Cfg::SMEM);
  const int m_blocks = (p.seqlen_q + Cfg::BM - 1) / Cfg::BM;
  dim3 grid(m_blocks * p.num_splits, p.h, p.b);
  kern<<<grid, 128, Cfg::SMEM, stream>>>(p);
  return cudaGetLastError();
}

template <typename T>
static cudaError_t launch_tile_d(const KParams& p, cudaStream_t stream) {
  switch (p.d) {
    case 32: return launch_tile_t<T, 32>(p, stream);
    case 64: return launch_tile_t<T, 64>(p, stream);
    case 96: return launch_tile_t<T, 96>(p, stream);
    case 128: return launch_tile_t<T, 128>(p, stream);
    case 160: return launch_tile_t<T, 160>(p, stream);
    case 192: return launch_tile_t<T, 192>(p, stream);
    case 224: return launch_tile_t<T, 224>(p, stream);
    case 256: return launch_tile_t<T, 256>(p, stream);
  }
  return cudaErrorInvalidValue;
}

bool tile_supported(int d) { return d >= 32 && d <= 256 && d % 32 == 0; }
int tile_block_m(int) { return 64; }

cudaError_t launch_tile(const KParams& p, bool is_bf16, cudaStream_t stream) {
  return is_bf16 ? launch_tile_d<__nv_bfloat16>(p, stream) : launch_tile_d<__half>(p, stream);
}

}  // namespace fa
