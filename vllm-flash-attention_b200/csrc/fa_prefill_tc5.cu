// fa_prefill_tc5.cu — compute-bound prefill / varlen kernel for sm_100a:
// TMA -> shared memory -> tcgen05.mma with S and O accumulators in TMEM -> online softmax in
// registers (one thread per query row, no shuffles) -> P back through TMEM -> PV MMA.
//
// Replaces compute_attn_1rowblock (reference src/flash_fwd_kernel.h:27-470; mma.sync m16n8k16 +
// cp.async, one 128-row tile per 4-warp CTA, single-buffered K/V).  The arithmetic contract is
// the reference's (src/softmax.h:66-186): fp32 scores, p = exp2(s*scale*log2e - m*scale*log2e),
// fp32 row sum of the unrounded p, P rounded to bf16/fp16 before P.V, LSE = m*scale + ln(sum).
// One deliberate difference: O is rescaled lazily (only when the running max grows by more than
// 2^8), which changes the reference max `m` used inside exp2 but not the normalised result.
//
// CTA = 256 query rows of one (sequence, head) as two 128-row tiles that ping-pong through the
// tensor core, 12 warps (warps 10-11 idle: setmaxnreg moves registers by whole warpgroups):
//   warps 0-3  softmax warpgroup of Q tile 0   (TMEM lanes 0..127 = rows; thread = row)
//   warps 4-7  softmax warpgroup of Q tile 1
//   warp  8    TMA producer (Q tiles once, then K_j, V_j through an NS-slot ring)
//   warp  9    MMA issuer (one elected lane) + TMEM allocation
// TMEM (512 columns): S0 [0,128) S1 [128,256) O0 [256,256+D) O1 [384,384+D); bf16 P_t overwrites
// the first 64 columns of S_t (all of S_t is in registers by then).
// MMA order:  QK0(0) QK1(0) { PV0(j) QK0(j+1) PV1(j) QK1(j+1) }  — while warpgroup 0 works on
// S0(j+1) the tensor core runs PV1(j) and QK1(j+1).
#include "fa_common.cuh"
#include "fa_sm100.cuh"
#include <type_traits>
#include <vector>
#include "fa_launch.h"

namespace fa {
using namespace sm100;

// Debug timeline (tools/tc5_trace.py): -DFA_TC5_TRACE builds record clock64 stamps of CTA (0,0,0):
// trace[role][iteration][event], role 0/1 = softmax warpgroup 0/1 (its first thread), 2 = MMA issuer.
#ifdef FA_TC5_TRACE
#define TC5_TRACE(role, iter, ev) do { if (p.trace && trace_cta && (iter) < 128) p.trace[((role) * 128 + (iter)) * 16 + (ev)] = clock64(); } while (0)
#else
#define TC5_TRACE(role, iter, ev) do { } while (0)
#endif

constexpr int TC5_BM = 128;        // rows per Q tile (UMMA M)
constexpr int TC5_BN = 128;        // keys per KV tile (UMMA N for QK^T)
constexpr int TC5_THREADS = 384;       // 3 warpgroups; warps 10-11 only donate registers
constexpr int TC5_REGS_SOFTMAX = 224, TC5_REGS_OTHER = 56;   // 256*224 + 128*56 = 384*168
constexpr int TC5_PRODUCER_WARP = 8, TC5_MMA_WARP = 9;

struct Tc5Maps { CUtensorMap q, k, v; };
template <typename T> struct IsBf16 { static constexpr bool value = false; };
template <> struct IsBf16<__nv_bfloat16> { static constexpr bool value = true; };

template <int D> struct Tc5Cfg {
  static constexpr int ND = D / 64;                         // 64-element (128 B) sub-tiles per row
  static constexpr int SUB_BYTES = 128 * 128;               // 128 rows x 128 B
  static constexpr int TILE_BYTES = ND * SUB_BYTES;         // a Q / K / V tile
  static constexpr int NS = D == 128 ? 4 : 8;
  static constexpr int OFF_Q = 0;
  static constexpr int OFF_KV = 2 * TILE_BYTES;
  static constexpr int OFF_BAR = OFF_KV + NS * TILE_BYTES;
  static constexpr int NUM_BARS = 2 /*q*/ + 2 * NS + 2 /*s*/ + 2 /*p*/ + 2 /*o*/;
  static constexpr int SMEM = OFF_BAR + NUM_BARS * 8 + 16 + 1024;   // +1024: manual alignment
  static constexpr int TM_S = 0, TM_O = 256;                // TMEM column bases; tile t adds t*128
};

// exp2 of 32 scaled scores -> 16 packed bf16/fp16 pairs + row-sum accumulation.
// The MUFU unit (16 ex2/clk/SM) needs exactly as many cycles per tile as the tensor pipe, so NEMU
// of every 16 pairs are evaluated on the FMA pipe instead (FA4-style): 2^x = 2^n * 2^r with
// n = round(x) taken from the magic-number add, r = x - n in [-0.5, 0.5], 2^r by a degree-3
// minimax polynomial (max rel. error 7.6e-5, 26x below the bf16 rounding of P that follows), and
// 2^n applied by adding n to the exponent field.  Everything but the exponent add is packed
// fp32x2.  Inputs are clamped at -126 (result 2^-126 instead of 0), so callers use NEMU = 0 on
// tiles that contain masked (-inf) scores.
#ifndef FA_TC5_EMU
#define FA_TC5_EMU 0
#endif
template <typename T, int NEMU>
__device__ __forceinline__ void tc5_exp_chunk(const uint32_t (&s)[32], const float2 mul2, const float2 nms2,
                                              float2& la, float2& lb, uint32_t (&pk)[16]) {
#pragma unroll
  for (int i = 0; i < 16; i++) {
    // packed fp32x2: one FFMA2 scales and shifts two scores, one FADD2 accumulates two sums
    float2 x = __ffma2_rn(make_float2(__uint_as_float(s[2 * i]), __uint_as_float(s[2 * i + 1])), mul2, nms2);
    float2 pp;
    if ((i & 1) && (i >> 1) < NEMU) {
      x.x = fmaxf(x.x, -126.f); x.y = fmaxf(x.y, -126.f);
      const float2 tt = __fadd2_rn(x, make_float2(12582912.f, 12582912.f));       // 1.5*2^23: low bits = round(x)
      const float2 fl = __fadd2_rn(tt, make_float2(-12582912.f, -12582912.f));
      const float2 r = __ffma2_rn(fl, make_float2(-1.f, -1.f), x);
      float2 q = __ffma2_rn(r, make_float2(0.05520550534f, 0.05520550534f), make_float2(0.2426139712f, 0.2426139712f));
      q = __ffma2_rn(q, r, make_float2(0.6932547688f, 0.6932547688f));
      q = __ffma2_rn(q, r, make_float2(0.9999276996f, 0.9999276996f));
      pp.x = __int_as_float(__float_as_int(q.x) + (__float_as_int(tt.x) << 23));
      pp.y = __int_as_float(__float_as_int(q.y) + (__float_as_int(tt.y) << 23));
    } else {
      pp = make_float2(fast_exp2(x.x), fast_exp2(x.y));
    }
    if (i & 1) lb = __fadd2_rn(lb, pp); else la = __fadd2_rn(la, pp);
    pk[i] = Pack2<T>::pack(pp.x, pp.y);
  }
}

// Per-Q-tile KV tile range [jb, je) (in units of TC5_BN keys); empty when the tile has no rows
// or no visible keys.
struct TileRange { int jb, je; };
__device__ __forceinline__ TileRange tc5_range(const KParams& p, int r0, int sq, int sk) {
  TileRange t; t.jb = 0; t.je = 0;
  if (r0 >= sq || sk <= 0) return t;
  int r1 = min(r0 + TC5_BM, sq) - 1;
  int lo = 0, hi = sk;
  if (p.is_causal) hi = min(sk, r1 + sk - sq + 1);
  else if (p.is_local) { hi = min(sk, r1 + sk - sq + 1 + p.window_right); lo = max(0, r0 + sk - sq - p.window_left); }
  if (hi <= lo) return t;
  t.jb = lo / TC5_BN; t.je = (hi + TC5_BN - 1) / TC5_BN;
  return t;
}

template <typename T, int D, bool HAS_ALIBI, bool PAGED>
__global__ void __launch_bounds__(TC5_THREADS, 1)
fa_prefill_tc5_kernel(const __grid_constant__ KParams p, const __grid_constant__ Tc5Maps maps) {
  using Cfg = Tc5Cfg<D>;
  constexpr int ND = Cfg::ND, NS = Cfg::NS;
  constexpr bool IS_BF16 = IsBf16<T>::value;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;   // SWIZZLE_128B needs 1024 B
  const uint32_t q_sm = smem_base + Cfg::OFF_Q, kv_sm = smem_base + Cfg::OFF_KV;
  const uint32_t bars = smem_base + Cfg::OFF_BAR;
  const uint32_t bar_q = bars;                    // [2]
  const uint32_t bar_kv_full = bars + 16;         // [NS]
  const uint32_t bar_kv_empty = bar_kv_full + 8 * NS;
  const uint32_t bar_s = bar_kv_empty + 8 * NS;   // [2] S_t ready
  const uint32_t bar_p = bar_s + 16;              // [2] P_t written (128 arrivals)
  const uint32_t bar_o = bar_p + 16;              // [2] PV_t(j) complete
  const uint32_t tmem_slot = bar_o + 16;
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int head = blockIdx.x, bidb = blockIdx.z;
  [[maybe_unused]] const bool trace_cta = blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && (threadIdx.x & 127) == 0 || (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && warp == TC5_MMA_WARP);
  const int m_block = p.is_causal ? (gridDim.y - 1 - blockIdx.y) : blockIdx.y;   // heavy tiles first
  const int kvh = head / p.group;
  const SeqInfo si(p, bidb);
  const int sq = si.sq, sk = si.sk;
  const int m0 = m_block * 2 * TC5_BM;
  if (m0 >= sq || sq * p.group < p.min_rows || sq * p.group > p.max_rows) return;
  const int q_row0 = p.cu_seqlens_q ? p.cu_seqlens_q[bidb] : 0;     // row coordinate bases (varlen)
  const int k_row0 = p.cu_seqlens_k ? p.cu_seqlens_k[bidb] : 0;
  const int bcoord = p.cu_seqlens_q ? 0 : bidb;

  const TileRange tr0 = tc5_range(p, m0, sq, sk), tr1 = tc5_range(p, m0 + TC5_BM, sq, sk);
  const bool act0 = tr0.je > tr0.jb, act1 = tr1.je > tr1.jb;
  const int j_begin = act0 ? (act1 ? min(tr0.jb, tr1.jb) : tr0.jb) : tr1.jb;
  const int j_end = act0 || act1 ? max(tr0.je, tr1.je) : j_begin;

  // ---- one-time setup ----
  if (warp == TC5_PRODUCER_WARP && lane == 0) {
    prefetch_tmap(&maps.q); prefetch_tmap(&maps.k); prefetch_tmap(&maps.v);
    mbar_init(bar_q, 1); mbar_init(bar_q + 8, 1);
    for (int i = 0; i < NS; i++) { mbar_init(bar_kv_full + 8 * i, 1); mbar_init(bar_kv_empty + 8 * i, 1); }
    for (int t = 0; t < 2; t++) { mbar_init(bar_s + 8 * t, 1); mbar_init(bar_p + 8 * t, 128); mbar_init(bar_o + 8 * t, 1); }
    fence_barrier_init();
  }
  if (warp == TC5_MMA_WARP) { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot_ptr;

  // register re-split by whole warpgroups (setmaxnreg sits at the top of each role so that ptxas
  // allocates each role against its own budget): the softmax threads hold a full 128-column S row
  if (warp == TC5_PRODUCER_WARP) {
    setmaxnreg_dec<TC5_REGS_OTHER>();
    // =============================== TMA producer ===============================
    if (elect_one()) {
      for (int t = 0; t < 2; t++) {
        if (!(t == 0 ? act0 : act1)) continue;
        mbar_expect_tx(bar_q + 8 * t, Cfg::TILE_BYTES);
        for (int c = 0; c < ND; c++)
          tma_load_4d(q_sm + t * Cfg::TILE_BYTES + c * Cfg::SUB_BYTES, &maps.q, bar_q + 8 * t,
                      c * 64, head, q_row0 + m0 + t * TC5_BM, bcoord);
      }
      if constexpr (!PAGED) {
        int item = 0;
        for (int j = j_begin; j < j_end; j++) {
          for (int kv = 0; kv < 2; kv++, item++) {
            const int slot = item % NS, round = item / NS;
            mbar_wait(bar_kv_empty + 8 * slot, (round & 1) ^ 1);
            mbar_expect_tx(bar_kv_full + 8 * slot, Cfg::TILE_BYTES);
            for (int c = 0; c < ND; c++)
              tma_load_4d(kv_sm + slot * Cfg::TILE_BYTES + c * Cfg::SUB_BYTES, kv == 0 ? &maps.k : &maps.v,
                          bar_kv_full + 8 * slot, c * 64, kvh, k_row0 + j * TC5_BN, p.cu_seqlens_k ? 0 : bidb);
          }
        }
      }
    }
    __syncwarp();
    if constexpr (PAGED) {
      // Paged K/V (reference src/utils.h:297-315, kernel_traits.h:136-144): a 128-key tile is
      // 128/box pages (or a slice of one big page); lane L looks up the block id of box L in the
      // block table and issues that box's TMA loads itself, so the table indirection costs one
      // coalesced load per tile.  The tensor maps are (d, heads, page_row, block).  The V tile that
      // contains the end of the sequence is special: rows >= seqlen_k hold whatever the pool holds
      // (possibly NaN) and 0 * NaN would poison P.V, so its boxes that reach past the end are copied
      // by hand (valid rows) / zero-filled (the rest) in the 128-B swizzle TMA would have used.
      const int page = p.page_block_size;
      const int box_rows = min(page, TC5_BN), nbox = TC5_BN / box_rows;
      const int* table = p.block_table + (int64_t)bidb * p.block_table_batch_stride;
      int item = 0;
      for (int j = j_begin; j < j_end; j++) {
        const int row0 = j * TC5_BN;
        const int valid = min(TC5_BN, sk - row0);                    // rows of this tile inside the sequence (>= 1)
        for (int kv = 0; kv < 2; kv++, item++) {
          const int slot = item % NS, round = item / NS;
          if (lane == 0) mbar_wait(bar_kv_empty + 8 * slot, (round & 1) ^ 1);
          __syncwarp();
          const uint32_t dst = kv_sm + slot * Cfg::TILE_BYTES, bar = bar_kv_full + 8 * slot;
          const bool tail = kv == 1 && valid < TC5_BN;
          const int n_tma_boxes = tail ? valid / box_rows : nbox;      // boxes entirely inside the sequence
          if (tail) {
            const T* vbase = reinterpret_cast<const T*>(p.v) + (int64_t)kvh * p.v_head_stride;
            const int r_begin = n_tma_boxes * box_rows;
            for (int idx = lane; idx < (TC5_BN - r_begin) * ND * 8; idx += 32) {
              const int ch = idx & 7, c = (idx >> 3) % ND, r = r_begin + (idx >> 3) / ND;
              uint4 val = make_uint4(0, 0, 0, 0);
              if (r < valid) {
                const int g = row0 + r, pg = g / page;
                const int blk = table[min(pg, p.max_blocks_per_seq - 1)];
                val = *reinterpret_cast<const uint4*>(vbase + (int64_t)blk * p.v_batch_stride +
                                                      (int64_t)(g - pg * page) * p.v_row_stride + c * 64 + ch * 8);
              }
              asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dst + c * Cfg::SUB_BYTES + r * 128 + ((ch ^ (r & 7)) << 4)),
                           "r"(val.x), "r"(val.y), "r"(val.z), "r"(val.w) : "memory");
            }
            fence_proxy_async_smem();
          }
          __syncwarp();
          if (lane == 0) mbar_expect_tx(bar, (uint32_t)(n_tma_boxes * box_rows * 128 * ND));
          __syncwarp();
          if (lane < n_tma_boxes) {
            const int g = row0 + lane * box_rows, pg = g / page;
            const int blk = table[min(pg, p.max_blocks_per_seq - 1)];
            for (int c = 0; c < ND; c++)
              tma_load_4d(dst + c * Cfg::SUB_BYTES + lane * box_rows * 128, kv == 0 ? &maps.k : &maps.v, bar,
                          c * 64, kvh, g - pg * page, blk);
          }
        }
      }
      __syncwarp();
    }
  } else if (warp == TC5_MMA_WARP) {
    setmaxnreg_dec<TC5_REGS_OTHER>();
    // =============================== MMA issuer ===============================
    // One thread issues both tiles' chains in the order PV0(j) QK0(j+1) PV1(j) QK1(j+1): the shared
    // in-order stream is what keeps the two softmax warpgroups in ANTI-phase (one exponentiates
    // while the other tile's MMAs run); two independent issuers drift into phase and halve the
    // MUFU rate each warpgroup sees.  Only one barrier wait (P_t) sits on a tile's chain: the
    // arrival of V_j and K_{j+1} is checked before it.
    constexpr uint32_t idesc_qk = make_idesc(IS_BF16, TC5_BM, TC5_BN, false);
    constexpr uint32_t idesc_pv = make_idesc(IS_BF16, TC5_BM, D, true);
    if (elect_one()) {
      auto active = [&](int t, int j) { return t == 0 ? (j >= tr0.jb && j < tr0.je) : (j >= tr1.jb && j < tr1.je); };
      // S_t = Q_t K_j^T : A = Q (K-major, SW128), B = K (K-major, SW128); D/16 k-steps
      auto issue_qk = [&](int t, int j, int slot, bool release) {
        const uint32_t qa = q_sm + t * Cfg::TILE_BYTES, kb = kv_sm + slot * Cfg::TILE_BYTES;
#pragma unroll
        for (int ks = 0; ks < D / 16; ks++) {
          const uint32_t off = (ks / 4) * Cfg::SUB_BYTES + (ks % 4) * 32;
          umma_ss(tmem + Cfg::TM_S + t * 128, make_desc_sw128(qa + off, 16, 1024),
                  make_desc_sw128(kb + off, 16, 1024), idesc_qk, ks > 0);
        }
        TC5_TRACE(2, j - 1 - j_begin, t * 8 + 5);
        umma_commit(bar_s + 8 * t);
        if (release) umma_commit(bar_kv_empty + 8 * slot);
      };
      if (act0) mbar_wait(bar_q, 0);
      if (act1) mbar_wait(bar_q + 8, 0);
      if (j_begin < j_end) {
        mbar_wait(bar_kv_full, 0);            // item 0 = K_{j_begin}: slot 0, first round
        tc_fence_after();
        const bool a0 = active(0, j_begin), a1 = active(1, j_begin);
        if (a0) issue_qk(0, j_begin, 0, !a1);
        if (a1) issue_qk(1, j_begin, 0, true);
      }
      for (int j = j_begin; j < j_end; j++) {
        const int vi = 2 * (j - j_begin) + 1, ki = vi + 1;
        const int vslot = vi % NS, kslot = ki % NS;
        const bool more = j + 1 < j_end;
        TC5_TRACE(2, j - j_begin, 0);
        mbar_wait(bar_kv_full + 8 * vslot, (vi / NS) & 1);
        if (more) mbar_wait(bar_kv_full + 8 * kslot, (ki / NS) & 1);
        TC5_TRACE(2, j - j_begin, 1);
        const bool a0 = active(0, j), a1 = active(1, j);
        const bool k0 = more && active(0, j + 1), k1 = more && active(1, j + 1);
#pragma unroll
        for (int t = 0; t < 2; t++) {
          const bool a = t == 0 ? a0 : a1, kq = t == 0 ? k0 : k1;
          const int jb_t = t == 0 ? tr0.jb : tr1.jb;
          if (a) {
            // O_t (+)= P_t V_j : A = P (TMEM), B = V (MN-major, SW128); 8 k-steps of 16 keys
            mbar_wait(bar_p + 8 * t, (uint32_t)((j - jb_t) & 1));
            TC5_TRACE(2, j - j_begin, t * 8 + 2);
            tc_fence_after();
            const uint32_t vb = kv_sm + vslot * Cfg::TILE_BYTES;
#pragma unroll
            for (int kk = 0; kk < TC5_BN / 16; kk++)
              umma_ts(tmem + Cfg::TM_O + t * 128, tmem + Cfg::TM_S + t * 128 + kk * 8,
                      make_desc_sw128(vb + kk * 2048, Cfg::SUB_BYTES, 1024), idesc_pv, (j > jb_t) || kk > 0);
            TC5_TRACE(2, j - j_begin, t * 8 + 3);
            umma_commit(bar_o + 8 * t);
            if (t == 1 || !a1) umma_commit(bar_kv_empty + 8 * vslot);
            TC5_TRACE(2, j - j_begin, t * 8 + 4);
          }
          if (kq) issue_qk(t, j + 1, kslot, t == 1 || !k1);
          TC5_TRACE(2, j - j_begin, t * 8 + 6);
        }
        if (!a0 && !a1) umma_commit(bar_kv_empty + 8 * vslot);          // unreachable for monotone windows
        if (more && !k0 && !k1) umma_commit(bar_kv_empty + 8 * kslot);
      }
    }
    __syncwarp();
  } else if (warp < 8) {
    setmaxnreg_inc<TC5_REGS_SOFTMAX>();
    // =============================== softmax warpgroups ===============================
#include "fa_prefill_tc5_softmax.inc"
  } else {
    setmaxnreg_dec<TC5_REGS_OTHER>();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == TC5_MMA_WARP) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

// =========================================================================================
// host side: tensor maps + launch
// =========================================================================================
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  }
  return fn;
}

// 4-D map over (d, heads, rows, batch) with a 64 x 1 x box_rows x 1 box and 128-byte swizzle.
static bool make_map(CUtensorMap* m, const void* base, bool bf16, int box_rows, int d, int heads, int64_t rows, int batch,
                     int64_t head_stride, int64_t row_stride, int64_t batch_stride) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return false;
  cuuint64_t dims[4] = {(cuuint64_t)d, (cuuint64_t)heads, (cuuint64_t)rows, (cuuint64_t)batch};
  cuuint64_t strides[3] = {(cuuint64_t)head_stride * 2, (cuuint64_t)row_stride * 2,
                           (cuuint64_t)(batch > 1 ? batch_stride : row_stride * rows) * 2};
  cuuint32_t box[4] = {64, 1, (cuuint32_t)box_rows, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(m, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4,
                  const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

bool tc5_supported(const KParams& p) {
  if (p.d != 64 && p.d != 128) return false;
  if (p.cache_batch_idx || p.seqused_k) return false;
  if (p.num_splits > 1) return false;
  if (p.block_table) {
    // paged K/V: a 128-key tile must be a whole number of pages, or a page a whole number of tiles
    const int pg = p.page_block_size;
    if (pg <= 0 || !(pg % TC5_BN == 0 || TC5_BN % pg == 0) || p.num_blocks <= 0) return false;
  } else {
    if (p.cu_seqlens_k && !p.is_seqlens_k_cumulative) return false;
    if ((p.cu_seqlens_q == nullptr) != (p.cu_seqlens_k == nullptr)) return false;
  }
  // TMA: 16-byte aligned base and strides
  auto al = [](const void* x) { return (reinterpret_cast<uintptr_t>(x) & 15) == 0; };
  if (!al(p.q) || !al(p.k) || !al(p.v) || !al(p.o)) return false;
  const int64_t st[] = {p.q_row_stride, p.q_head_stride, p.k_row_stride, p.k_head_stride, p.v_row_stride,
                        p.v_head_stride, p.o_row_stride, p.o_head_stride, p.q_batch_stride, p.k_batch_stride,
                        p.v_batch_stride, p.o_batch_stride};
  for (int64_t s : st) if (s % 8 != 0 || s < 0) return false;
  return get_encode_fn() != nullptr;
}

template <typename T, int D, bool HAS_ALIBI, bool PAGED>
static cudaError_t launch_tc5_t(const KParams& p, const Tc5Maps& maps, cudaStream_t stream) {
  using Cfg = Tc5Cfg<D>;
  auto kern = fa_prefill_tc5_kernel<T, D, HAS_ALIBI, PAGED>;
  FA_SET_SMEM_ONCE(kern, Cfg::SMEM);
  dim3 grid(p.h, (p.seqlen_q + 2 * TC5_BM - 1) / (2 * TC5_BM), p.b);
  kern<<<grid, TC5_THREADS, Cfg::SMEM, stream>>>(p, maps);
  return cudaGetLastError();
}

template <typename T, int D>
static cudaError_t launch_tc5_d(const KParams& p, const Tc5Maps& maps, cudaStream_t stream) {
  const bool alibi = p.alibi_slopes != nullptr;
  if (p.block_table) return alibi ? launch_tc5_t<T, D, true, true>(p, maps, stream) : launch_tc5_t<T, D, false, true>(p, maps, stream);
  return alibi ? launch_tc5_t<T, D, true, false>(p, maps, stream) : launch_tc5_t<T, D, false, false>(p, maps, stream);
}

#ifdef FA_TC5_TRACE
// Debug only: a device buffer for the timeline, dumped to $FA_B200_TRACE_FILE after each launch.
static cudaError_t launch_prefill_tc5_impl(const KParams& p, bool is_bf16, int64_t total_q, int64_t total_k, cudaStream_t stream);
cudaError_t launch_prefill_tc5(const KParams& p_in, bool is_bf16, int64_t total_q, int64_t total_k, cudaStream_t stream) {
  static long long* buf = nullptr;
  const size_t n = 3 * 128 * 16;
  if (!buf) cudaMalloc(&buf, n * sizeof(long long));
  cudaMemsetAsync(buf, 0, n * sizeof(long long), stream);
  KParams p = p_in; p.trace = buf;
  cudaError_t e = launch_prefill_tc5_impl(p, is_bf16, total_q, total_k, stream);
  const char* path = getenv("FA_B200_TRACE_FILE");
  if (e == cudaSuccess && path) {
    std::vector<long long> h(n);
    cudaStreamSynchronize(stream);
    cudaMemcpy(h.data(), buf, n * sizeof(long long), cudaMemcpyDeviceToHost);
    FILE* f = fopen(path, "wb"); if (f) { fwrite(h.data(), sizeof(long long), n, f); fclose(f); }
  }
  return e;
}
static cudaError_t launch_prefill_tc5_impl(const KParams& p, bool is_bf16, int64_t total_q, int64_t total_k, cudaStream_t stream) {
#else
cudaError_t launch_prefill_tc5(const KParams& p, bool is_bf16, int64_t total_q, int64_t total_k, cudaStream_t stream) {
#endif
  Tc5Maps maps;
  const bool varlen = p.cu_seqlens_q != nullptr;
  const int64_t q_rows = varlen ? total_q : p.seqlen_q;
  bool ok = make_map(&maps.q, p.q, is_bf16, TC5_BM, p.d, p.h, q_rows, varlen ? 1 : p.b, p.q_head_stride, p.q_row_stride, p.q_batch_stride);
  if (p.block_table) {   // (d, heads, row in page, block)
    const int box = p.page_block_size < TC5_BN ? p.page_block_size : TC5_BN;
    ok = ok && make_map(&maps.k, p.k, is_bf16, box, p.d, p.h_k, p.page_block_size, p.num_blocks, p.k_head_stride, p.k_row_stride, p.k_batch_stride) &&
         make_map(&maps.v, p.v, is_bf16, box, p.d, p.h_k, p.page_block_size, p.num_blocks, p.v_head_stride, p.v_row_stride, p.v_batch_stride);
  } else {
    const bool kvarlen = p.cu_seqlens_k != nullptr;
    const int64_t k_rows = kvarlen ? total_k : p.seqlen_k;
    ok = ok && make_map(&maps.k, p.k, is_bf16, TC5_BN, p.d, p.h_k, k_rows, kvarlen ? 1 : p.b, p.k_head_stride, p.k_row_stride, p.k_batch_stride) &&
         make_map(&maps.v, p.v, is_bf16, TC5_BN, p.d, p.h_k, k_rows, kvarlen ? 1 : p.b, p.v_head_stride, p.v_row_stride, p.v_batch_stride);
  }
  if (!ok) return cudaErrorInvalidValue;
  if (is_bf16) return p.d == 64 ? launch_tc5_d<__nv_bfloat16, 64>(p, maps, stream) : launch_tc5_d<__nv_bfloat16, 128>(p, maps, stream);
  return p.d == 64 ? launch_tc5_d<__half, 64>(p, maps, stream) : launch_tc5_d<__half, 128>(p, maps, stream);
}

}  // namespace fa
