// fa_prefill_tc5.cu — compute-bound prefill / varlen kernel for sm_100a:
// TMA -> shared memory -> tcgen05.mma with S and O accumulators in TMEM -> online softmax in
// registers (one thread per query row, no shuffles) -> P back through TMEM (or smem) -> PV MMA.
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
#include "fa_launch.h"

namespace fa {
using namespace sm100;

constexpr int TC5_BM = 128;        // rows per Q tile (UMMA M)
constexpr int TC5_BN = 128;        // keys per KV tile (UMMA N for QK^T)
constexpr int TC5_THREADS = 384;       // 3 warpgroups; warps 10-11 only donate registers
constexpr int TC5_REGS_SOFTMAX = 224, TC5_REGS_OTHER = 56;   // 256*224 + 128*56 = 384*168
constexpr int TC5_PRODUCER_WARP = 8, TC5_MMA_WARP = 9;

struct Tc5Maps { CUtensorMap q, k, v; };
template <typename T> struct IsBf16 { static constexpr bool value = false; };
template <> struct IsBf16<__nv_bfloat16> { static constexpr bool value = true; };

template <int D, bool P_TMEM> struct Tc5Cfg {
  static constexpr int ND = D / 64;                         // 64-element (128 B) sub-tiles per row
  static constexpr int SUB_BYTES = 128 * 128;               // 128 rows x 128 B
  static constexpr int TILE_BYTES = ND * SUB_BYTES;         // a Q / K / V tile
  static constexpr int P_BYTES = P_TMEM ? 0 : 2 * SUB_BYTES;  // 128 x 128 bf16 per Q tile
  static constexpr int NS = P_TMEM ? (D == 128 ? 5 : 8) : (D == 128 ? 3 : 6);
  static constexpr int OFF_Q = 0;
  static constexpr int OFF_KV = 2 * TILE_BYTES;
  static constexpr int OFF_P = OFF_KV + NS * TILE_BYTES;
  static constexpr int OFF_BAR = OFF_P + 2 * P_BYTES;
  static constexpr int NUM_BARS = 2 /*q*/ + 2 * NS + 2 /*s*/ + 2 /*p*/ + 2 /*o*/;
  static constexpr int SMEM = OFF_BAR + NUM_BARS * 8 + 16 + 1024;   // +1024: manual alignment
  static constexpr int TM_S = 0, TM_O = 256;                // TMEM column bases; tile t adds t*128
};

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

template <typename T, int D, bool P_TMEM, bool HAS_ALIBI>
__global__ void __launch_bounds__(TC5_THREADS, 1)
fa_prefill_tc5_kernel(const __grid_constant__ KParams p, const __grid_constant__ Tc5Maps maps) {
  using Cfg = Tc5Cfg<D, P_TMEM>;
  constexpr int ND = Cfg::ND, NS = Cfg::NS;
  constexpr bool IS_BF16 = IsBf16<T>::value;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;   // SWIZZLE_128B needs 1024 B
  const uint32_t q_sm = smem_base + Cfg::OFF_Q, kv_sm = smem_base + Cfg::OFF_KV, p_sm = smem_base + Cfg::OFF_P;
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
  const int m_block = p.is_causal ? (gridDim.y - 1 - blockIdx.y) : blockIdx.y;   // heavy tiles first
  const int kvh = head / p.group;
  const SeqInfo si(p, bidb);
  const int sq = si.sq, sk = si.sk;
  const int m0 = m_block * 2 * TC5_BM;
  if (m0 >= sq) return;
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
    __syncwarp();
  } else if (warp == TC5_MMA_WARP) {
    setmaxnreg_dec<TC5_REGS_OTHER>();
    // =============================== MMA issuer ===============================
    constexpr uint32_t idesc_qk = make_idesc(IS_BF16, TC5_BM, TC5_BN, false);
    constexpr uint32_t idesc_pv = make_idesc(IS_BF16, TC5_BM, D, true);
    if (elect_one()) {
      auto active = [&](int t, int j) { return t == 0 ? (j >= tr0.jb && j < tr0.je) : (j >= tr1.jb && j < tr1.je); };
      auto slot_of = [&](int item, uint32_t& par) { par = (item / NS) & 1; return item % NS; };
      // S_t = Q_t K_j^T : A = Q (K-major, SW128), B = K (K-major, SW128); D/16 k-steps
      auto issue_qk = [&](int t, int j) {
        uint32_t par; const int slot = slot_of(2 * (j - j_begin), par);
        mbar_wait(bar_kv_full + 8 * slot, par);
        tc_fence_after();
        const uint32_t qa = q_sm + t * Cfg::TILE_BYTES, kb = kv_sm + slot * Cfg::TILE_BYTES;
#pragma unroll
        for (int ks = 0; ks < D / 16; ks++) {
          const uint32_t off = (ks / 4) * Cfg::SUB_BYTES + (ks % 4) * 32;
          umma_ss(tmem + Cfg::TM_S + t * 128, make_desc_sw128(qa + off, 16, 1024),
                  make_desc_sw128(kb + off, 16, 1024), idesc_qk, ks > 0);
        }
        umma_commit(bar_s + 8 * t);
        const bool last_user = t == 1 || !active(1, j);
        if (last_user) umma_commit(bar_kv_empty + 8 * slot);
      };
      // O_t (+)= P_t V_j : A = P (TMEM, or smem K-major), B = V (MN-major, SW128); 8 k-steps of 16 keys
      auto issue_pv = [&](int t, int j, int jb_t, uint32_t p_par) {
        uint32_t par; const int slot = slot_of(2 * (j - j_begin) + 1, par);
        mbar_wait(bar_kv_full + 8 * slot, par);
        mbar_wait(bar_p + 8 * t, p_par);
        tc_fence_after();
        const uint32_t vb = kv_sm + slot * Cfg::TILE_BYTES;
#pragma unroll
        for (int kk = 0; kk < TC5_BN / 16; kk++) {
          const uint64_t bdesc = make_desc_sw128(vb + kk * 2048, Cfg::SUB_BYTES, 1024);
          if constexpr (P_TMEM) {
            umma_ts(tmem + Cfg::TM_O + t * 128, tmem + Cfg::TM_S + t * 128 + kk * 8, bdesc, idesc_pv, (j > jb_t) || kk > 0);
          } else {
            const uint32_t pa = p_sm + t * Cfg::P_BYTES + (kk / 4) * Cfg::SUB_BYTES + (kk % 4) * 32;
            umma_ss(tmem + Cfg::TM_O + t * 128, make_desc_sw128(pa, 16, 1024), bdesc, idesc_pv, (j > jb_t) || kk > 0);
          }
        }
        umma_commit(bar_o + 8 * t);
        const bool last_user = t == 1 || !active(1, j);
        if (last_user) umma_commit(bar_kv_empty + 8 * slot);
      };
      if (act0) mbar_wait(bar_q, 0);
      if (act1) mbar_wait(bar_q + 8, 0);
      for (int t = 0; t < 2; t++) if (active(t, j_begin)) issue_qk(t, j_begin);
      for (int j = j_begin; j < j_end; j++) {
        for (int t = 0; t < 2; t++) {
          const int jb_t = t == 0 ? tr0.jb : tr1.jb;
          if (active(t, j)) issue_pv(t, j, jb_t, (uint32_t)((j - jb_t) & 1));
          if (j + 1 < j_end && active(t, j + 1)) issue_qk(t, j + 1);
        }
        if (!active(0, j) && !active(1, j)) {   // unreachable for monotone windows; keeps the ring moving
          uint32_t par; int slot = slot_of(2 * (j - j_begin) + 1, par);
          mbar_wait(bar_kv_full + 8 * slot, par); umma_commit(bar_kv_empty + 8 * slot);
        }
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

// 4-D map over (d, heads, rows, batch) with a 64 x 1 x 128 x 1 box and 128-byte swizzle.
static bool make_map(CUtensorMap* m, const void* base, bool bf16, int d, int heads, int64_t rows, int batch,
                     int64_t head_stride, int64_t row_stride, int64_t batch_stride) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return false;
  cuuint64_t dims[4] = {(cuuint64_t)d, (cuuint64_t)heads, (cuuint64_t)rows, (cuuint64_t)batch};
  cuuint64_t strides[3] = {(cuuint64_t)head_stride * 2, (cuuint64_t)row_stride * 2,
                           (cuuint64_t)(batch > 1 ? batch_stride : row_stride * rows) * 2};
  cuuint32_t box[4] = {64, 1, 128, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(m, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4,
                  const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

static int g_tc5_p_mode = -1;   // -1: env not read; 0 = P through shared memory, 1 = P through TMEM
static int tc5_p_mode() {
  if (g_tc5_p_mode < 0) {
    const char* e = getenv("FA_B200_TC5_P");
    g_tc5_p_mode = (e && e[0] == 's') ? 0 : 1;
  }
  return g_tc5_p_mode;
}

bool tc5_supported(const KParams& p) {
  if (p.d != 64 && p.d != 128) return false;
  if (p.block_table || p.cache_batch_idx || p.seqused_k) return false;
  if (p.cu_seqlens_k && !p.is_seqlens_k_cumulative) return false;
  if ((p.cu_seqlens_q == nullptr) != (p.cu_seqlens_k == nullptr)) return false;
  if (p.num_splits > 1) return false;
  // TMA: 16-byte aligned base and strides
  auto al = [](const void* x) { return (reinterpret_cast<uintptr_t>(x) & 15) == 0; };
  if (!al(p.q) || !al(p.k) || !al(p.v) || !al(p.o)) return false;
  const int64_t st[] = {p.q_row_stride, p.q_head_stride, p.k_row_stride, p.k_head_stride, p.v_row_stride,
                        p.v_head_stride, p.o_row_stride, p.o_head_stride, p.q_batch_stride, p.k_batch_stride,
                        p.v_batch_stride, p.o_batch_stride};
  for (int64_t s : st) if (s % 8 != 0 || s < 0) return false;
  return get_encode_fn() != nullptr;
}

template <typename T, int D, bool P_TMEM, bool HAS_ALIBI>
static cudaError_t launch_tc5_t(const KParams& p, const Tc5Maps& maps, cudaStream_t stream) {
  using Cfg = Tc5Cfg<D, P_TMEM>;
  auto kern = fa_prefill_tc5_kernel<T, D, P_TMEM, HAS_ALIBI>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM);
  if (e != cudaSuccess) return e;
  dim3 grid(p.h, (p.seqlen_q + 2 * TC5_BM - 1) / (2 * TC5_BM), p.b);
  kern<<<grid, TC5_THREADS, Cfg::SMEM, stream>>>(p, maps);
  return cudaGetLastError();
}

template <typename T, int D>
static cudaError_t launch_tc5_d(const KParams& p, const Tc5Maps& maps, cudaStream_t stream) {
  const bool alibi = p.alibi_slopes != nullptr;
  if (tc5_p_mode() == 1) {
    return alibi ? launch_tc5_t<T, D, true, true>(p, maps, stream) : launch_tc5_t<T, D, true, false>(p, maps, stream);
  } else {
    return alibi ? launch_tc5_t<T, D, false, true>(p, maps, stream) : launch_tc5_t<T, D, false, false>(p, maps, stream);
  }
}

cudaError_t launch_prefill_tc5(const KParams& p, bool is_bf16, int64_t total_q, int64_t total_k, cudaStream_t stream) {
  Tc5Maps maps;
  const bool varlen = p.cu_seqlens_q != nullptr;
  const int64_t q_rows = varlen ? total_q : p.seqlen_q, k_rows = varlen ? total_k : p.seqlen_k;
  const int nb = varlen ? 1 : p.b;
  if (!make_map(&maps.q, p.q, is_bf16, p.d, p.h, q_rows, nb, p.q_head_stride, p.q_row_stride, p.q_batch_stride) ||
      !make_map(&maps.k, p.k, is_bf16, p.d, p.h_k, k_rows, nb, p.k_head_stride, p.k_row_stride, p.k_batch_stride) ||
      !make_map(&maps.v, p.v, is_bf16, p.d, p.h_k, k_rows, nb, p.v_head_stride, p.v_row_stride, p.v_batch_stride))
    return cudaErrorInvalidValue;
  if (is_bf16) return p.d == 64 ? launch_tc5_d<__nv_bfloat16, 64>(p, maps, stream) : launch_tc5_d<__nv_bfloat16, 128>(p, maps, stream);
  return p.d == 64 ? launch_tc5_d<__half, 64>(p, maps, stream) : launch_tc5_d<__half, 128>(p, maps, stream);
}

}  // namespace fa
