// fa_decode.cu — GQA-packed, paged, split-KV decode kernel (HBM-bound).
//
// Replaces compute_attn_1rowblock_splitkv (reference src/flash_fwd_kernel.h:474-1053) for the
// decode shape: few query rows (seqlen_q * group <= 16*MT) against a long (paged) KV cache.
//
// Design (not the reference's): one CTA per (split, kv_head, sequence).  All `group` query heads
// that share the KV head (times seqlen_q positions) are packed as the M rows of one m16 MMA tile,
// so K/V are streamed from HBM exactly once per KV head (the reference gets this through the
// seqlenq_ngroups_swapped reshape, flash_api.cpp:298-304).  The KV range of the CTA is cut into
// 16-row chunks (one page when page_block_size==16; never straddles a page since page%16==0,
// flash_api.cpp:491,735).  Chunks are dealt round-robin to the warps; EVERY WARP RUNS ITS OWN
// cp.async ring (STAGES deep) and its own online softmax over its chunks — no block-wide barrier
// in the streaming loop — and the NW per-warp (m,l,O) states are merged once at the end in a
// fixed order (bitwise deterministic).  The page id of chunk i+1 is fetched while chunk i is
// being issued, so the block_table indirection (reference src/utils.h:297-315) is off the
// critical path.  Loads are 16 B cp.async, 2 full 256 B rows per warp instruction.
#include "fa_common.cuh"
#include "fa_warp_attn.cuh"
#include "fa_launch.h"
#include "fa_rotary.cuh"

namespace fa {

template <typename T, int D, int MT, int NW, int STAGES>
struct DecodeCfg {
  static constexpr int ROWS = 16 * MT;
  static constexpr int CHUNK = 16;
  static constexpr int TILE_BYTES = CHUNK * D * 2;             // one K (or V) chunk
  static constexpr int Q_BYTES = ROWS * D * 2;
  static constexpr int RING_BYTES = NW * STAGES * 2 * TILE_BYTES;
  static constexpr int RED_STRIDE = D + 8;                      // floats
  static constexpr int RED_BYTES = ROWS * RED_STRIDE * 4;
  static constexpr int STAT_BYTES = (2 * NW + 2) * ROWS * 4;    // m[NW][ROWS], l[NW][ROWS], inv_l, lse
  static constexpr int MERGE_BYTES = RED_BYTES + ROWS * 128 * 4;   // fused tail: O rows + the (row, split) weights of up to 128 splits
  static constexpr int MAIN_BYTES = RING_BYTES > MERGE_BYTES ? RING_BYTES : MERGE_BYTES;
  static constexpr int SMEM = Q_BYTES + MAIN_BYTES + STAT_BYTES;
};

// One piece of one (sequence, kv head) unit: chunks [c_begin, c_end) of its KV range, as piece `split` of `nsp`.
// cb_given < 0: the classic grid — piece `split` of p.num_splits equal pieces, range computed here; otherwise the range is
// handed in (balanced-range kernel below).
template <typename T, int D, int MT, int NW, int STAGES>
__device__ __forceinline__ void decode_unit(const KParams& p, uint8_t* smem, const int split, const int kvh, const int bidb,
                                            const int nsp, const int cb_given, const int ce_given, const int slot0 = 0) {
  using Cfg = DecodeCfg<T, D, MT, NW, STAGES>;
  constexpr int ROWS = Cfg::ROWS;
  constexpr int CPR = D / 8;  // 16 B chunks per row
  uint8_t* q_sm = smem;
  uint8_t* ring = smem + Cfg::Q_BYTES;
  float* stat_m = reinterpret_cast<float*>(smem + Cfg::Q_BYTES + Cfg::MAIN_BYTES);
  float* stat_l = stat_m + NW * ROWS;
  float* row_inv = stat_l + NW * ROWS;
  float* row_lse = row_inv + ROWS;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = p.group;
  const SeqInfo si(p, bidb);
  const int sq = si.sq, sk = si.sk;
  const int R = sq * g;  // live packed rows: r = qpos*g + head_in_group
  if (sq <= 0 || R < p.min_rows || R > p.max_rows) return;   // mixed batch: longer sequences belong to the prefill kernel

  // ---- KV chunk range of this CTA ----
  const RowWindow w_first = row_window(p, 0, sq, sk), w_last = row_window(p, sq - 1, sq, sk);
  const int c_lo = max(w_first.lo, 0) / 16;
  const int c_hi = (max(w_last.hi, 0) + 15) / 16;
  const int n_chunks = max(c_hi - c_lo, 0);
  const int per_split = (n_chunks + nsp - 1) / nsp;
  const int c_begin = cb_given >= 0 ? cb_given : c_lo + split * per_split;
  const int c_end = cb_given >= 0 ? min(ce_given, c_hi) : min(c_hi, c_begin + per_split);

  // ---- in-kernel append (flash_fwd_kernel.h:672-824): the call's new K/V rows, rotated, go into the cache here instead
  // of in a launch of their own.  Row L + i (L = length before the call) is written by the ONE CTA of this KV head whose
  // chunk range holds it — one CTA per (sequence, kv head, split), so no two CTAs write the same row (the reference
  // documents a benign race between the query heads of a group, :678-680) — made visible device-wide, and only then do
  // this CTA's warps start streaming chunks; every other CTA never reads those rows.
  if (p.knew != nullptr) {
    const int L = sk - p.seqlen_knew;
    const int r_lo = max(L, c_begin * 16), r_hi = min(sk, c_end * 16);
    const int half = p.rotary_dim / 2;
    for (int pos = r_lo + warp; pos < r_hi; pos += NW) {
      const int i = pos - L;
      int64_t koff, voff;
      if (p.block_table) {
        const int pg = min(pos / p.page_block_size, p.max_blocks_per_seq - 1);
        const int blk = p.block_table[(int64_t)bidb * p.block_table_batch_stride + pg];
        const int in_page = pos - (pos / p.page_block_size) * p.page_block_size;
        koff = (int64_t)blk * p.k_batch_stride + (int64_t)in_page * p.k_row_stride;
        voff = (int64_t)blk * p.v_batch_stride + (int64_t)in_page * p.v_row_stride;
      } else {
        if (pos >= p.seqlen_k) continue;          // cache full: the reference would write out of bounds
        koff = si.k_off + (int64_t)pos * p.k_row_stride;
        voff = si.v_off + (int64_t)pos * p.v_row_stride;
      }
      const T* ksrc = reinterpret_cast<const T*>(p.knew) + (int64_t)bidb * p.knew_batch_stride + (int64_t)i * p.knew_row_stride + (int64_t)kvh * p.knew_head_stride;
      const T* vsrc = reinterpret_cast<const T*>(p.vnew) + (int64_t)bidb * p.vnew_batch_stride + (int64_t)i * p.vnew_row_stride + (int64_t)kvh * p.vnew_head_stride;
      T* kdst = const_cast<T*>(reinterpret_cast<const T*>(p.k)) + koff + (int64_t)kvh * p.k_head_stride;
      T* vdst = const_cast<T*>(reinterpret_cast<const T*>(p.v)) + voff + (int64_t)kvh * p.v_head_stride;
      const int rp = min(pos, max(p.seqlen_ro - 1, 0));
      rotate_row<T>(ksrc, kdst, D, p.rotary_cos ? p.rotary_dim : 0, p.rotary_interleaved != 0,
                    reinterpret_cast<const T*>(p.rotary_cos) + (int64_t)rp * half,
                    reinterpret_cast<const T*>(p.rotary_sin) + (int64_t)rp * half, lane, 32);
      rotate_row<T>(vsrc, vdst, D, 0, true, nullptr, nullptr, lane, 32);
    }
    if (r_lo < r_hi) __threadfence();
    __syncthreads();
  }

  // ---- Q: gather the packed rows into shared memory (zero rows beyond R) ----
  if (p.rotary_cos != nullptr) {
    // rotary on q (rotary.h:19-145): row qpos at position L (+ qpos under a causal / local mask), rotated on its way
    // into the tile — the rotated q never exists in global memory
    const T* qb = reinterpret_cast<const T*>(p.q) + si.q_off;
    const int L = sk - p.seqlen_knew, half = p.rotary_dim / 2;
    for (int r = warp; r < ROWS; r += NW) {
      if (r < R) {
        const int qpos = r / g, hg = r - qpos * g;
        const int rp = min(L + (p.rot_by_qpos ? qpos : 0), max(p.seqlen_ro - 1, 0));
        const T* src = qb + (int64_t)qpos * p.q_row_stride + (int64_t)(kvh * g + hg) * p.q_head_stride;
        rotate_row_to<T>(src, [&](int c) { return reinterpret_cast<T*>(q_sm + swz<D>(r, c)); }, D, p.rotary_dim, p.rotary_interleaved != 0,
                         reinterpret_cast<const T*>(p.rotary_cos) + (int64_t)rp * half,
                         reinterpret_cast<const T*>(p.rotary_sin) + (int64_t)rp * half, lane, 32);
      } else {
        for (int c = lane; c < CPR; c += 32) *reinterpret_cast<uint4*>(q_sm + swz<D>(r, c)) = make_uint4(0, 0, 0, 0);
      }
    }
    cp_async_commit();      // (empty) keeps the group count of the waits below
  } else {
    const T* qb = reinterpret_cast<const T*>(p.q) + si.q_off;
    for (int idx = threadIdx.x; idx < ROWS * CPR; idx += NW * 32) {
      int r = idx / CPR, c = idx % CPR;
      int qpos = r / g, hg = r - qpos * g;
      const T* src = qb + (int64_t)qpos * p.q_row_stride + (int64_t)(kvh * g + hg) * p.q_head_stride + c * 8;
      cp_async16(smem_u32(q_sm) + swz<D>(r, c), r < R ? (const void*)src : (const void*)qb, r < R);
    }
    cp_async_commit();
  }

  // ---- per-thread row metadata (window, ALiBi) for its 2*MT rows ----
  int row_lo[MT][2], row_hi[MT][2], row_shift[MT][2];
  float row_slope[MT][2];
  const bool has_alibi = p.alibi_slopes != nullptr;
#pragma unroll
  for (int i = 0; i < MT; i++)
#pragma unroll
    for (int r = 0; r < 2; r++) {
      int row = i * 16 + (lane >> 2) + r * 8;
      int qpos = min(row / g, sq - 1), hg = row % g;
      RowWindow w = row_window(p, qpos, sq, sk);
      row_lo[i][r] = w.lo; row_hi[i][r] = w.hi; row_shift[i][r] = qpos + sk - sq;
      row_slope[i][r] = has_alibi
          ? p.alibi_slopes[bidb * p.alibi_batch_stride + kvh * g + hg] * 1.4426950408889634f : 0.f;
    }
  const int hi_min = w_first.hi, lo_max = w_last.lo;

  // ---- K/V addressing ----
  const T* kbase = reinterpret_cast<const T*>(p.k) + (int64_t)kvh * p.k_head_stride;
  const T* vbase = reinterpret_cast<const T*>(p.v) + (int64_t)kvh * p.v_head_stride;
  const bool paged = p.block_table != nullptr;
  const int* table = paged ? p.block_table + (int64_t)bidb * p.block_table_batch_stride : nullptr;
  if (!paged) { kbase += si.k_off; vbase += si.v_off; }
  const int page = p.page_block_size;
  auto page_of = [&](int c) -> int {  // physical block id of chunk c (clamped: SURVEY 5a.8)
    int pg = min((c * 16) / page, p.max_blocks_per_seq - 1);
    return table[pg];
  };

  uint8_t* my_ring = ring + warp * (STAGES * 2 * Cfg::TILE_BYTES);
  auto issue = [&](int c, int stage, int blk) {
    const int row0 = c * 16;
    int64_t koff, voff;
    if (paged) {
      int in_page = row0 % page;
      koff = (int64_t)blk * p.k_batch_stride + (int64_t)in_page * p.k_row_stride;
      voff = (int64_t)blk * p.v_batch_stride + (int64_t)in_page * p.v_row_stride;
    } else {
      koff = (int64_t)row0 * p.k_row_stride;
      voff = (int64_t)row0 * p.v_row_stride;
    }
    uint32_t ks = smem_u32(my_ring + stage * 2 * Cfg::TILE_BYTES);
    uint32_t vs = ks + Cfg::TILE_BYTES;
#pragma unroll
    for (int it = 0; it < (16 * CPR) / 32; it++) {
      int seg = it * 32 + lane;
      int r = seg / CPR, cc = seg % CPR;
      bool ok = row0 + r < sk;
      int rr = ok ? r : 0;
      cp_async16(ks + swz<D>(r, cc), kbase + koff + (int64_t)rr * p.k_row_stride + cc * 8, ok);
      cp_async16(vs + swz<D>(r, cc), vbase + voff + (int64_t)rr * p.v_row_stride + cc * 8, ok);
    }
  };

  WarpAttn<T, D, MT, 16> attn;
  attn.init();

  const int first = c_begin + warp;
  const int n_mine = first < c_end ? (c_end - first + NW - 1) / NW : 0;

  // prologue: STAGES-1 chunks in flight
  int blk_next = 0;
#pragma unroll
  for (int s = 0; s < STAGES - 1; s++) {
    if (s < n_mine) {
      int c = first + s * NW;
      issue(c, s, paged ? page_of(c) : 0);
    }
    cp_async_commit();
  }
  if (paged && STAGES - 1 < n_mine) blk_next = page_of(first + (STAGES - 1) * NW);

  const uint32_t q_s = smem_u32(q_sm);
  // Q visible to all warps
  cp_async_wait<STAGES - 1>();   // Q group is the oldest
  __syncthreads();

  for (int i = 0; i < n_mine; i++) {
    const int c = first + i * NW;
    {  // refill the stage consumed in the previous iteration
      int inext = i + STAGES - 1;
      if (inext < n_mine) {
        issue(first + inext * NW, inext % STAGES, blk_next);
        if (paged && inext + 1 < n_mine) blk_next = page_of(first + (inext + 1) * NW);
      }
      cp_async_commit();
    }
    cp_async_wait<STAGES - 1>();
    __syncwarp();
    const int stage = i % STAGES;
    const uint32_t ks = smem_u32(my_ring + stage * 2 * Cfg::TILE_BYTES);
    const int col0 = c * 16;
    const bool need_mask = has_alibi || (col0 + 16 > hi_min) || (col0 < lo_max);
    attn.step(q_s, ks, ks + Cfg::TILE_BYTES, p.scale_log2, need_mask,
              [&](int row, int col, float x) -> float {
                const int i2 = row >> 4, r2 = (row >> 3) & 1;
                const int j = col0 + col;
                if (has_alibi) x += p.is_causal ? row_slope[i2][r2] * (float)j     // src/mask.h:181
                                                : -row_slope[i2][r2] * fabsf((float)(row_shift[i2][r2] - j));
                return (j >= row_lo[i2][r2] && j < row_hi[i2][r2]) ? x : -INFINITY;
              });
    __syncwarp();
  }
  attn.finish_sums();
  cp_async_wait<0>();

  // ---- merge the NW warp states (fixed order => deterministic) ----
  if ((lane & 3) == 0) {
#pragma unroll
    for (int i = 0; i < MT; i++)
#pragma unroll
      for (int r = 0; r < 2; r++) {
        int row = i * 16 + (lane >> 2) + r * 8;
        stat_m[warp * ROWS + row] = attn.m[i][r];
        stat_l[warp * ROWS + row] = attn.l[i][r];
      }
  }
  __syncthreads();   // also: every warp is done with its ring -> reuse as the reduction buffer
  float my_scale[MT][2];
#pragma unroll
  for (int i = 0; i < MT; i++)
#pragma unroll
    for (int r = 0; r < 2; r++) {
      int row = i * 16 + (lane >> 2) + r * 8;
      float mt = -INFINITY;
#pragma unroll
      for (int w = 0; w < NW; w++) mt = fmaxf(mt, stat_m[w * ROWS + row]);
      float lt = 0.f;
#pragma unroll
      for (int w = 0; w < NW; w++) {
        float mw = stat_m[w * ROWS + row];
        lt += mw == -INFINITY ? 0.f : stat_l[w * ROWS + row] * fast_exp2(mw - mt);
      }
      my_scale[i][r] = attn.m[i][r] == -INFINITY ? 0.f : fast_exp2(attn.m[i][r] - mt);
      if (warp == 0 && (lane & 3) == 0) {
        const bool empty = !(lt > 0.f);
        row_inv[row] = empty ? 0.f : 1.f / lt;
        // natural-log LSE = max*scale + ln(sum) (softmax.h:179); m is in log2 units
        row_lse[row] = empty ? ((nsp > 1 || p.partial || p.split_counters != nullptr || p.xchg_world > 0) ? -INFINITY : INFINITY)
                             : mt * 0.6931471805599453f + __logf(lt);
      }
    }
  float* red = reinterpret_cast<float*>(ring);
  for (int w = 0; w < NW; w++) {
    if (warp == w) {
#pragma unroll
      for (int i = 0; i < MT; i++)
#pragma unroll
        for (int j = 0; j < D / 8; j++)
#pragma unroll
          for (int r = 0; r < 2; r++) {
            int row = i * 16 + (lane >> 2) + r * 8;
            float2* dst = reinterpret_cast<float2*>(red + row * Cfg::RED_STRIDE + j * 8 + (lane & 3) * 2);
            float2 val = make_float2(attn.o[i][j][2 * r] * my_scale[i][r], attn.o[i][j][2 * r + 1] * my_scale[i][r]);
            if (w > 0) { float2 old = *dst; val.x += old.x; val.y += old.y; }
            *dst = val;
          }
    }
    __syncthreads();
  }

  // ---- write out ----
  const int Rw = min(R, ROWS);
  if (p.split_counters != nullptr || p.xchg_world > 0) {
    // ================= fused tail: split merge (+ cross-GPU exchange) inside this kernel =================
    // Replaces the separate fa_combine_kernel launch (combine_attn_seqk_parallel, flash_fwd_kernel.h:1092-1250) and,
    // for sequence-sharded decode, the separate exchange kernel: the LAST CTA to finish a (sequence, kv head) unit
    // merges the unit's split partials, and — sharded — pushes the merged rows to every rank over NVLink, waits
    // for the peers' rows of the same unit and writes the final result.  Per-unit state only: no CTA waits for
    // another CTA of this grid, so nothing has to be co-resident.
    __shared__ int s_last;
    __shared__ unsigned s_n;
    const int tid = threadIdx.x;
    const int unit = bidb * p.h_k + kvh;
    auto head_of = [&](int r, int& qpos) { qpos = r / g; return kvh * g + (r - qpos * g); };
    // where piece s of this unit keeps row r of its partial: row index prow(r) + s * sstride of o_accum / lse_accum.
    // Classic grid: the reference's (split, b, h, sq) layout.  Balanced ranges: a POOL of piece slots of p.sk_rows rows,
    // the unit's pieces in consecutive slots from slot0 (= first CTA of the unit + unit index: unique, and at most
    // gridDim.x + b*h_k slots exist however the ranges fall).
    const bool pool = cb_given >= 0;
    const int64_t sstride = pool ? (int64_t)p.sk_rows : (int64_t)p.b * p.h * p.part_seq_stride;
    auto prow = [&](int r) -> int64_t {
      if (pool) return (int64_t)slot0 * p.sk_rows + r;
      int qpos; const int hd = head_of(r, qpos);
      return ((int64_t)bidb * p.h + hd) * p.part_seq_stride + qpos;
    };
    if (nsp > 1) {
      // a. this split's partial, fp32 (split, b, h, sq[, d]) — the layout of the reference's accum buffers
      for (int idx = tid; idx < Rw * (D / 4); idx += NW * 32) {
        const int r = idx / (D / 4), c4 = idx % (D / 4);
        const int64_t orow = prow(r) + split * sstride;
        const float inv = row_inv[r];
        float4 x = *reinterpret_cast<const float4*>(red + r * Cfg::RED_STRIDE + c4 * 4);
        x.x *= inv; x.y *= inv; x.z *= inv; x.w *= inv;
        *reinterpret_cast<float4*>(p.o_accum + orow * D + c4 * 4) = x;
      }
      for (int r = tid; r < Rw; r += NW * 32) p.lse_accum[prow(r) + split * sstride] = row_lse[r];
      // b. arrival ticket: the last CTA of the unit carries on, the others are done
      __threadfence();
      __syncthreads();
      if (tid == 0) {
        const int t = atomicAdd(p.split_counters + unit, 1);
        s_last = t == nsp - 1;
        if (s_last) p.split_counters[unit] = 0;          // ready for the next launch
      }
      __syncthreads();
      if (!s_last) return;
      __threadfence();
      // c. merge the unit's splits into red / row_lse.  First the weights exp(lse_s - lse) of every (row, split) into
      // shared memory (one warp per row, lanes own splits); then every thread owns (row, 4-column) items and streams the
      // splits with 16 independent 16-byte loads in flight — the partials sit in L2, and a walk with one or two loads in
      // flight per lane is what made this tail cost more than the streaming loop of a short split.  Fixed order:
      // deterministic.
      const int S = nsp;
      float* wsm = red + ROWS * Cfg::RED_STRIDE;            // [Rw][S] weights, behind the O rows (the ring is free by now)
      for (int r = warp; r < Rw; r += NW) {
        const int64_t row0 = prow(r);
        float ls[4], mx = -INFINITY;
#pragma unroll
        for (int i = 0; i < 4; i++) {
          const int s2 = lane + 32 * i;
          ls[i] = s2 < S ? __ldcg(p.lse_accum + s2 * sstride + row0) : -INFINITY;
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
        const float lse = empty ? -INFINITY : logf(sum) + mx_safe;
#pragma unroll
        for (int i = 0; i < 4; i++) {
          const int s2 = lane + 32 * i;
          if (s2 < S) wsm[r * S + s2] = empty ? 0.f : expf(ls[i] - lse);
        }
        if (lane == 0) row_lse[r] = lse;
      }
      __syncthreads();
      for (int idx = tid; idx < Rw * (D / 4); idx += NW * 32) {
        const int r = idx / (D / 4), c4 = idx % (D / 4);
        const float* src = p.o_accum + prow(r) * D + c4 * 4;
        const float* wr = wsm + r * S;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        int s2 = 0;
        for (; s2 + 16 <= S; s2 += 16) {
          float4 x[16];
#pragma unroll
          for (int u = 0; u < 16; u++) x[u] = __ldcg(reinterpret_cast<const float4*>(src + (s2 + u) * sstride * D));
#pragma unroll
          for (int u = 0; u < 16; u++) {
            const float w = wr[s2 + u];
            acc.x += w * x[u].x; acc.y += w * x[u].y; acc.z += w * x[u].z; acc.w += w * x[u].w;
          }
        }
        for (; s2 + 8 <= S; s2 += 8) {
          float4 x[8];
#pragma unroll
          for (int u = 0; u < 8; u++) x[u] = __ldcg(reinterpret_cast<const float4*>(src + (s2 + u) * sstride * D));
#pragma unroll
          for (int u = 0; u < 8; u++) {
            const float w = wr[s2 + u];
            acc.x += w * x[u].x; acc.y += w * x[u].y; acc.z += w * x[u].z; acc.w += w * x[u].w;
          }
        }
        for (; s2 < S; s2++) {
          const float4 x = __ldcg(reinterpret_cast<const float4*>(src + s2 * sstride * D));
          const float w = wr[s2];
          acc.x += w * x.x; acc.y += w * x.y; acc.z += w * x.z; acc.w += w * x.w;
        }
        *reinterpret_cast<float4*>(red + r * Cfg::RED_STRIDE + c4 * 4) = acc;
      }
      __syncthreads();
    } else {
      // one split: this CTA's own result is the unit's; normalise it in place (row_lse already holds -inf for empty rows)
      for (int idx = tid; idx < Rw * (D / 4); idx += NW * 32) {
        const int r = idx / (D / 4), c4 = idx % (D / 4);
        const float inv = row_inv[r];
        float4 x = *reinterpret_cast<const float4*>(red + r * Cfg::RED_STRIDE + c4 * 4);
        x.x *= inv; x.y *= inv; x.z *= inv; x.w *= inv;
        *reinterpret_cast<float4*>(red + r * Cfg::RED_STRIDE + c4 * 4) = x;
      }
      __syncthreads();
    }
    // red[r][0..D) = this rank's normalised fp32 O rows of the unit, row_lse[r] = natural-log LSE (-inf: no key)

    if (p.xchg_world > 0) {
      // d. exchange: call number n of THIS unit; slot (n & 1, source rank, unit) of every rank's receive area is
      // private to the unit, so units whose call numbers differ (varying batch sizes) never touch each other's data;
      // a rank cannot run two calls ahead of a peer for the same unit (its call n+1 needs the peer's lines of call
      // n+1, sent after the peer finished call n).  The rows travel as FLAGGED LINES (fa_common.cuh: two fp32, each
      // in one 64-bit word with n): the data is its own arrival flag, so a rank's rows cost the peers one one-way trip —
      // no system-scope release fence (a round trip over NVSwitch) and no separate flag behind it.
      if (tid == 0) s_n = p.xchg_state[unit] + 1u;
      __syncthreads();
      const unsigned n = s_n;
      const int W = p.xchg_world;
      constexpr int LPR = D / 2;                                     // lines per row
      const int o_lines = Rw * LPR, all_lines = o_lines + (Rw + 1) / 2;   // O rows, then the LSEs two per line
      // slot base in floats: a slot holds unit_elems values = unit_elems 64-bit words
      auto slot = [&](int src) { return ((int64_t)((n & 1u) * W + src) * p.xchg_max_units + unit) * p.xchg_unit_elems * 2; };
      const int64_t mine = slot(p.xchg_rank);
      for (int idx = tid; idx < all_lines; idx += NW * 32) {
        float a, b;
        if (idx < o_lines) {
          const int r = idx / LPR, c2 = idx - r * LPR;
          const float2 x = *reinterpret_cast<const float2*>(red + r * Cfg::RED_STRIDE + 2 * c2);
          a = x.x; b = x.y;
        } else {
          const int r = 2 * (idx - o_lines);
          a = row_lse[r]; b = r + 1 < Rw ? row_lse[r + 1] : 0.f;
        }
        for (int w = 0; w < W; w++)
          if (w != p.xchg_rank) st_flagged(p.xchg_recv[w] + mine + (int64_t)idx * 4, a, b, n);
      }
      __syncthreads();                                                 // every thread has read red / row_lse for the push
      // e. merge the ranks' rows as they arrive: one warp per row; own rows straight from shared memory, fixed rank
      // order (every rank computes the same sums)
      const float* base = p.xchg_recv[p.xchg_rank];
      for (int r = warp; r < Rw; r += NW) {
        float my_ls = -INFINITY;
        if (lane < W) {
          if (lane == p.xchg_rank) my_ls = row_lse[r];
          else { const float2 t = ld_flagged(base + slot(lane) + (int64_t)(o_lines + (r >> 1)) * 4, n); my_ls = (r & 1) ? t.y : t.x; }
        }
        float ls[FA_KP_MAX_PEERS], mx = -INFINITY;
#pragma unroll
        for (int w = 0; w < FA_KP_MAX_PEERS; w++) {
          ls[w] = __shfl_sync(0xffffffffu, my_ls, w);                  // lanes >= W hold -inf
          mx = fmaxf(mx, ls[w]);
        }
        const bool empty = mx == -INFINITY;
        float sum = 0.f;
#pragma unroll
        for (int w = 0; w < FA_KP_MAX_PEERS; w++) sum += empty ? 0.f : expf(ls[w] - mx);
        const float lse = empty ? -INFINITY : logf(sum) + mx;
        float wg[FA_KP_MAX_PEERS];
#pragma unroll
        for (int w = 0; w < FA_KP_MAX_PEERS; w++) wg[w] = empty ? 0.f : expf(ls[w] - lse);
        for (int c2 = lane; c2 < LPR; c2 += 32) {
          float2 acc = make_float2(0.f, 0.f);
          float2* own = reinterpret_cast<float2*>(red + r * Cfg::RED_STRIDE + 2 * c2);
#pragma unroll
          for (int w = 0; w < FA_KP_MAX_PEERS; w++) {
            if (w < W) {
              const float2 x = w == p.xchg_rank ? *own : ld_flagged(base + slot(w) + (int64_t)(r * LPR + c2) * 4, n);
              acc.x += wg[w] * x.x; acc.y += wg[w] * x.y;
            }
          }
          *own = acc;
        }
        __syncwarp();
        if (lane == 0) row_lse[r] = lse;
      }
      if (tid == 0) p.xchg_state[unit] = n;
      __syncthreads();
    }

    // f. final rows: fp32 record (partial-result mode) or the output dtype; LSE +inf for rows without any key
    if (p.partial) {
      for (int idx = tid; idx < Rw * (D / 4); idx += NW * 32) {
        const int r = idx / (D / 4), c4 = idx % (D / 4);
        int qpos; const int hd = head_of(r, qpos);
        const int64_t orow = ((int64_t)bidb * p.h + hd) * p.lse_seq_stride + qpos;
        *reinterpret_cast<float4*>(reinterpret_cast<float*>(p.o) + orow * D + c4 * 4) = *reinterpret_cast<const float4*>(red + r * Cfg::RED_STRIDE + c4 * 4);
      }
      for (int r = tid; r < Rw; r += NW * 32) {
        int qpos; const int hd = head_of(r, qpos);
        p.lse[((int64_t)bidb * p.h + hd) * p.lse_seq_stride + qpos] = row_lse[r];
      }
    } else {
      T* ob = reinterpret_cast<T*>(p.o) + o_seq_offset(p, bidb);
      for (int idx = tid; idx < Rw * CPR; idx += NW * 32) {
        const int r = idx / CPR, c = idx % CPR;
        int qpos; const int hd = head_of(r, qpos);
        const float* src = red + r * Cfg::RED_STRIDE + c * 8;
        const float4 a = *reinterpret_cast<const float4*>(src), b2 = *reinterpret_cast<const float4*>(src + 4);
        uint4 pk;
        pk.x = Pack2<T>::pack(a.x, a.y); pk.y = Pack2<T>::pack(a.z, a.w);
        pk.z = Pack2<T>::pack(b2.x, b2.y); pk.w = Pack2<T>::pack(b2.z, b2.w);
        *reinterpret_cast<uint4*>(ob + (int64_t)qpos * p.o_row_stride + (int64_t)hd * p.o_head_stride + c * 8) = pk;
      }
      if (p.lse) {
        for (int r = tid; r < Rw; r += NW * 32) {
          int qpos; const int hd = head_of(r, qpos);
          const float l = row_lse[r];
          p.lse[((int64_t)bidb * p.h + hd) * p.lse_seq_stride + qpos] = l == -INFINITY ? INFINITY : l;
        }
      }
    }
    return;
  }
  if ((nsp > 1 || p.partial)) {
    // fp32 partials, (split, b, h, sq[, d]) like the reference's accum buffers (flash_api.cpp:212-213)
    for (int idx = threadIdx.x; idx < Rw * (D / 4); idx += NW * 32) {
      int r = idx / (D / 4), c4 = idx % (D / 4);
      int qpos = r / g, hg = r - qpos * g;
      int64_t orow = (((int64_t)split * p.b + bidb) * p.h + kvh * g + hg) * p.lse_seq_stride + qpos;
      float inv = row_inv[r];
      float4 x = *reinterpret_cast<const float4*>(red + r * Cfg::RED_STRIDE + c4 * 4);
      x.x *= inv; x.y *= inv; x.z *= inv; x.w *= inv;
      *reinterpret_cast<float4*>(p.o_accum + orow * D + c4 * 4) = x;
    }
    for (int r = threadIdx.x; r < Rw; r += NW * 32) {
      int qpos = r / g, hg = r - qpos * g;
      int64_t orow = (((int64_t)split * p.b + bidb) * p.h + kvh * g + hg) * p.lse_seq_stride + qpos;
      p.lse_accum[orow] = row_lse[r];
    }
  } else {
    T* ob = reinterpret_cast<T*>(p.o) + o_seq_offset(p, bidb);
    for (int idx = threadIdx.x; idx < Rw * CPR; idx += NW * 32) {
      int r = idx / CPR, c = idx % CPR;
      int qpos = r / g, hg = r - qpos * g;
      float inv = row_inv[r];
      const float* src = red + r * Cfg::RED_STRIDE + c * 8;
      float4 a = *reinterpret_cast<const float4*>(src), b2 = *reinterpret_cast<const float4*>(src + 4);
      uint4 pk;
      pk.x = Pack2<T>::pack(a.x * inv, a.y * inv); pk.y = Pack2<T>::pack(a.z * inv, a.w * inv);
      pk.z = Pack2<T>::pack(b2.x * inv, b2.y * inv); pk.w = Pack2<T>::pack(b2.z * inv, b2.w * inv);
      *reinterpret_cast<uint4*>(ob + (int64_t)qpos * p.o_row_stride + (int64_t)(kvh * g + hg) * p.o_head_stride + c * 8) = pk;
    }
    if (p.lse) {
      for (int r = threadIdx.x; r < Rw; r += NW * 32) {
        int qpos = r / g, hg = r - qpos * g;
        p.lse[((int64_t)bidb * p.h + kvh * g + hg) * p.lse_seq_stride + qpos] = row_lse[r];
      }
    }
  }
}

template <typename T, int D, int MT, int NW, int STAGES>
__global__ void __launch_bounds__(NW * 32) fa_decode_kernel(const __grid_constant__ KParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  decode_unit<T, D, MT, NW, STAGES>(p, smem, blockIdx.x, blockIdx.y, blockIdx.z, p.num_splits, -1, -1);
}

// Balanced ranges ("stream-K" over the KV axis) for launches with FEW units (b * h_k below a few waves of CTAs) — small
// decode batches, TP shards with one or two KV heads, the decode half of a mixed batch.  One CTA per (unit, equal piece)
// then suffers twice: pieces of ragged sequences differ in size, and piece counts quantise badly against the SM count
// (192 units x 3 pieces = 1.95 waves of unequal CTAs: 3.8 TB/s where the same bytes in equal shares stream at 4.6).  Here
// the grid is a fixed 2 CTAs per SM and the chunks of ALL units, laid end to end, are cut into gridDim.x equal ranges:
// CTA c owns chunks [c * per, (c + 1) * per) of that line, i.e. the tail of one unit, maybe whole units, the head of
// another — each such segment is one call of decode_unit with its piece index and the unit's piece count, and the unit's
// last arriver merges the pieces as before (same fused tail, same arrival counters, same cross-GPU exchange).  Every CTA
// derives the same layout from the sequence lengths itself (one pass over the batch, a prefix sum in shared memory):
// no scheduling kernel in front.  `per` is at least p.sk_min_per_cta, which bounds the pieces of a unit and so the
// partial buffers (host: fa_api.cu).
constexpr int SK_MAX_B = 512;
template <typename T, int D, int MT, int NW, int STAGES>
__global__ void __launch_bounds__(NW * 32) fa_decode_streamk_kernel(const __grid_constant__ KParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ int s_w[SK_MAX_B], s_clo[SK_MAX_B], s_pref[SK_MAX_B + 1];
  __shared__ int s_wmax, s_units;
  const int tid = threadIdx.x;
  for (int b = tid; b < p.b; b += NW * 32) {
    const SeqInfo si(p, b);
    const int R = si.sq * p.group;
    int w = 0, clo = 0;
    if (si.sq > 0 && R >= p.min_rows && R <= p.max_rows) {            // served by this launch (mixed batches: decode sequences only)
      const RowWindow wf = row_window(p, 0, si.sq, si.sk), wl = row_window(p, si.sq - 1, si.sq, si.sk);
      clo = max(wf.lo, 0) / 16;
      w = max(max((max(wl.hi, 0) + 15) / 16 - clo, 0), 1);            // a sequence without visible keys still has its rows written
    }
    s_w[b] = w; s_clo[b] = clo;
  }
  __syncthreads();
  if (tid < 32) {                                                      // exclusive prefix sum over <= 512 sequences: one warp, 16 per lane
    int loc[SK_MAX_B / 32], sum = 0;
#pragma unroll
    for (int i = 0; i < SK_MAX_B / 32; i++) { const int b = tid * (SK_MAX_B / 32) + i; loc[i] = sum; sum += b < p.b ? s_w[b] : 0; }
    int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, incl, o); if (tid >= o) incl += y; }
    const int base = incl - sum;
#pragma unroll
    for (int i = 0; i < SK_MAX_B / 32; i++) { const int b = tid * (SK_MAX_B / 32) + i; if (b < p.b) s_pref[b] = base + loc[i]; }
    int wmax = 0;
#pragma unroll
    for (int i = 0; i < SK_MAX_B / 32; i++) { const int b = tid * (SK_MAX_B / 32) + i; if (b < p.b) wmax = max(wmax, s_w[b]); }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, o));
    int served = 0;
#pragma unroll
    for (int i = 0; i < SK_MAX_B / 32; i++) { const int b = tid * (SK_MAX_B / 32) + i; if (b < p.b && s_w[b] > 0) served++; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) served += __shfl_xor_sync(0xffffffffu, served, o);
    if (tid == 31) { s_pref[p.b] = incl; s_wmax = wmax; s_units = served; }
  }
  __syncthreads();
  const long long total = (long long)s_pref[p.b] * p.h_k;
  // chunks per CTA: an equal share; at least p.sk_min_per_cta (a warp should see a few chunks); and enough that no unit
  // is cut into more pieces than the merge takes (128)
  // (the floor must not pile whole units onto one CTA while others idle: at least one CTA per served unit)
  const long long n_act = min((long long)gridDim.x, max((long long)s_units * p.h_k, (total + p.sk_min_per_cta - 1) / p.sk_min_per_cta));
  const int per = max((int)((total + n_act - 1) / max(n_act, 1ll)), (s_wmax + 126) / 127);
  long long g = (long long)blockIdx.x * per;
  const long long g_end = min(total, g + per);
  while (g < g_end) {
    // sequence that holds global chunk g: largest b with h_k * pref[b] <= g (sequences with weight 0 are skipped by the search)
    int lo = 0, hi = p.b;
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if ((long long)s_pref[mid] * p.h_k <= g) lo = mid; else hi = mid; }
    const int b = lo, w = s_w[b];
    const long long in_seq = g - (long long)s_pref[b] * p.h_k;
    const int kvh = (int)(in_seq / w), local = (int)(in_seq - (long long)kvh * w);
    const int seg = (int)min((long long)(w - local), g_end - g);
    const long long unit0 = g - local;                                 // first global chunk of this unit
    const int first_cta = (int)(unit0 / per), last_cta = (int)((unit0 + w - 1) / per);
    decode_unit<T, D, MT, NW, STAGES>(p, smem, (int)blockIdx.x - first_cta, kvh, b, last_cta - first_cta + 1,
                                      s_clo[b] + local, s_clo[b] + local + seg, first_cta + b * p.h_k + kvh);
    g += seg;
    __syncthreads();                                                   // shared memory is reused by the next segment
  }
}

template <typename T, int D, int MT, int NW, int STAGES>
static cudaError_t launch_decode_t(const KParams& p, cudaStream_t stream) {
  using Cfg = DecodeCfg<T, D, MT, NW, STAGES>;
  if (p.sk_grid > 0) {
    auto kern = fa_decode_streamk_kernel<T, D, MT, NW, STAGES>;
    FA_SET_SMEM_ONCE(kern, Cfg::SMEM);
    kern<<<p.sk_grid, NW * 32, Cfg::SMEM, stream>>>(p);
    return cudaGetLastError();
  }
  auto kern = fa_decode_kernel<T, D, MT, NW, STAGES>;
  FA_SET_SMEM_ONCE(kern, Cfg::SMEM);
  dim3 grid(p.num_splits, p.h_k, p.b);
  kern<<<grid, NW * 32, Cfg::SMEM, stream>>>(p);
  return cudaGetLastError();
}

template <typename T>
static cudaError_t launch_decode_d(const KParams& p, int mt, cudaStream_t stream) {
  // ring depth: about 24 KB of K+V in flight per warp
// small head dims move few bytes per 16-key chunk, so they need more warps per SM to cover the per-chunk
// latency chain (measured on B200, batch 256 x ctx 4096: d=64 4.97 -> 6.25 TB/s, d=32 3.48 -> 4.41 TB/s going
// from 4 to 8 warps per CTA at the same shared-memory footprint)
#ifndef FA_DEC32_NW
#define FA_DEC32_NW 8
#define FA_DEC32_ST 4
#endif
#ifndef FA_DEC64_NW
#define FA_DEC64_NW 8
#define FA_DEC64_ST 3
#endif
#ifndef FA_DEC96_NW
#define FA_DEC96_NW 8
#define FA_DEC96_ST 2
#endif
#ifndef FA_DEC160_ST
#define FA_DEC160_ST 2      // 2 stages: two CTAs per SM instead of one (d=160 4.72 -> 6.68 TB/s, d=192 5.83 -> 6.82 TB/s)
#define FA_DEC192_ST 2
#endif
#ifndef FA_DEC224_NW
#define FA_DEC224_NW 6      // one CTA per SM either way: 6 warps x 2 stages beat 4 x 3 (d=224 5.98 -> 6.46, d=256 6.64 -> 6.91 TB/s)
#define FA_DEC224_ST 2
#define FA_DEC256_NW 6
#define FA_DEC256_ST 2
#endif
#ifndef FA_DEC128_NW
#define FA_DEC128_NW 4
#define FA_DEC128_ST 3
#endif
#define FA_DEC_CASE(D_, NW_, ST_)                                                    \
  case D_:                                                                           \
    if (mt == 1) return launch_decode_t<T, D_, 1, NW_, ST_>(p, stream);              \
    if constexpr (D_ <= 128) { if (mt == 2) return launch_decode_t<T, D_, 2, NW_, ST_>(p, stream); } \
    break;
  // d = 128, all pieces resident at once and long (fa_api.cu sets wide_cta): 12 warps per SM keep more loads in flight than
  // 8 — 6 warps x 2 stages, same ring bytes.  C5 on one GPU: B=8 0.623 -> 0.589 ms, B=1 0.0939 -> 0.0916 ms; contexts of
  // 16k+ 2-4 % (profiles/r02_decode_wide_cta.txt).  Batches of several waves (C3) stay at 4 x 3: 6.83 against 6.74 TB/s.
  // Up to 16 packed rows only: the 32-row form needs 245 registers per thread, and 192 such threads leave one CTA per SM.
  if (p.d == 128 && p.wide_cta == 1 && p.sk_grid == 0 && mt == 1) return launch_decode_t<T, 128, 1, 6, 2>(p, stream);
  // Five waves and more (wide_cta == 2): 4 warps x 2 stages, a 64 KB ring, THREE CTAs per SM — a launch's fixed cost is paid
  // per wave of resident CTAs (zero-length contexts, 2048 units: 32.8 -> 24.6 us) and 12 warps per SM hold up better in
  // the sustained regime (2048 units x 4096 keys, 200 replays: 680 -> 654 us; the bench's 20-call C3 is the same either way).
  if (p.d == 128 && p.wide_cta == 2 && p.sk_grid == 0 && mt == 1) return launch_decode_t<T, 128, 1, 4, 2>(p, stream);
  switch (p.d) {
    FA_DEC_CASE(32, FA_DEC32_NW, FA_DEC32_ST)
    FA_DEC_CASE(64, FA_DEC64_NW, FA_DEC64_ST)
    FA_DEC_CASE(96, FA_DEC96_NW, FA_DEC96_ST)
    FA_DEC_CASE(128, FA_DEC128_NW, FA_DEC128_ST)
    FA_DEC_CASE(160, 4, FA_DEC160_ST)
    FA_DEC_CASE(192, 4, FA_DEC192_ST)
    FA_DEC_CASE(224, FA_DEC224_NW, FA_DEC224_ST)
    FA_DEC_CASE(256, FA_DEC256_NW, FA_DEC256_ST)
    default: break;
  }
#undef FA_DEC_CASE
  return cudaErrorInvalidValue;
}

// Every head dimension of the reference build (multiples of 32 up to 256, static_switch.h:80-107):
// up to 32 packed rows while the O fragment fits the register file (d <= 128), 16 above.
bool decode_supported(int d, int rows) {
  if (d < 32 || d > 256 || d % 32 != 0) return false;
  return rows <= (d <= 128 ? 32 : 16);
}

cudaError_t launch_decode(const KParams& p, bool is_bf16, cudaStream_t stream) {
  int rows = p.seqlen_q * p.group;
  if (rows > p.max_rows) rows = p.max_rows;   // mixed batch: longer sequences are skipped by the kernel
  int mt = rows <= 16 ? 1 : 2;
  return is_bf16 ? launch_decode_d<__nv_bfloat16>(p, mt, stream) : launch_decode_d<__half>(p, mt, stream);
}

}  // namespace fa
