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
// Persistent CTAs (one per SM) draw work units from a cost-sorted list (tc5_schedule_kernel).
// Work unit = 256 query rows of one (sequence, head) as two 128-row tiles that ping-pong through the
// tensor core, 12 warps (warps 10-11 idle: setmaxnreg moves registers by whole warpgroups):
//   warps 0-3  softmax warpgroup of Q tile 0   (TMEM lanes 0..127 = rows; thread = row)
//   warps 4-7  softmax warpgroup of Q tile 1
//   warp  8    scheduler + TMA producer (work mailbox, Q tiles, then K_j, V_j through an NS-slot ring)
//   warp  9    MMA issuer (one elected lane) + TMEM allocation
// TMEM (512 columns): S0 [0,128) S1 [128,256) O0 [256,256+D) O1 [384,384+D); bf16 P_t overwrites
// the first 64 columns of S_t (all of S_t is in registers by then).
// MMA order:  QK0(0) QK1(0) { PV0(j) QK0(j+1) PV1(j) QK1(j+1) }  — while warpgroup 0 works on
// S0(j+1) the tensor core runs PV1(j) and QK1(j+1).
// Head dims: instantiations D = 64 / 128 / 256 serve d = 32,64 / 96,128 / 160..256 (d is a run-time value:
// TMA boxes clip at the true width, QK^T takes d/16 k-steps, PV has N = d).  D = 256 runs ONE Q tile per
// work unit (TMEM: S [0,128) Q [128,256) O [256,512); two softmax threads per row, fa_prefill_tc5_softmax_pair.inc;
// Q, K, V share a 3-slot ring of 64 KB tiles) — see Tc5Cfg::TILES.
#include "fa_common.cuh"
#include "fa_sm100.cuh"
#include <algorithm>
#include <cstring>
#include <type_traits>
#include <vector>
#include "fa_launch.h"

namespace fa {
using namespace sm100;

// Debug timeline (tools/tc5_trace.py): -DFA_TC5_TRACE builds record clock64 stamps of CTA (0,0,0):
// trace[role][iteration][event], role 0/1 = softmax warpgroup 0/1 (its first thread), 2 = MMA issuer.
#ifdef FA_TC5_TRACE
#define TC5_TRACE(role, iter, ev) do { if (p.trace && trace_cta && (iter) < 127) p.trace[((role) * 128 + (iter)) * 16 + (ev)] = clock64(); } while (0)
#else
#define TC5_TRACE(role, iter, ev) do { } while (0)
#endif

constexpr int TC5_BM = 128;        // rows per Q tile (UMMA M)
constexpr int TC5_BN = 128;        // keys per KV tile (UMMA N for QK^T)
constexpr int TC5_THREADS = 384;       // 3 warpgroups; warps 10-11 idle (setmaxnreg works on whole warpgroups)
// Registers live in the 4 SM sub-partitions (16K each): a sub-partition holds softmax warps w and w+4 plus
// one of warps 8-11, so 2*224 + 56 (= 3*168, the launch allocation) is the most the split can give.
constexpr int TC5_REGS_SOFTMAX = 224, TC5_REGS_OTHER = 56;
constexpr int TC5_PRODUCER_WARP = 8, TC5_MMA_WARP = 9;
#ifndef FA_TC5_GOVERNOR
#define FA_TC5_GOVERNOR 0
#endif
#ifndef FA_TC5_GOV_MARGIN
#define FA_TC5_GOV_MARGIN 24
#endif
__device__ __forceinline__ uint32_t clock32() { uint32_t c; asm volatile("mov.u32 %0, %%clock;" : "=r"(c)); return c; }

struct Tc5Maps { CUtensorMap q, k, v, o; };   // o: 64 x 1 x 32 x 1 box (one warp's epilogue slab)
template <typename T> struct IsBf16 { static constexpr bool value = false; };
template <> struct IsBf16<__nv_bfloat16> { static constexpr bool value = true; };

template <int D> struct Tc5Cfg {
  static constexpr int ND = D / 64;                         // 64-element (128 B) sub-tiles per row
  static constexpr int SUB_BYTES = 128 * 128;               // 128 rows x 128 B
  static constexpr int TILE_BYTES = ND * SUB_BYTES;         // a Q / K / V tile
  // Q tiles per work unit.  d <= 128: two 128-row tiles ping-pong through the tensor core.  128 < d <= 256
  // (D = 256): O alone takes 256 TMEM columns, so a unit is ONE 128-row tile that both softmax warpgroups
  // share (two threads per row).
  static constexpr int TILES = D > 128 ? 1 : 2;
  static constexpr int UNIT_ROWS = TILES * 128;
  // one-tile units: Q is only staged in shared memory on its way to TMEM, so it travels through the K/V ring
  // as one more item (order per unit: Q, K_jb, V_jb, K_jb+1, ...) and the ring is 3 slots of 64 KB
  static constexpr int NS = D == 256 ? 3 : (D == 128 ? 4 : 8);
  static constexpr int OFF_Q = 0;
  static constexpr int OFF_KV = TILES == 1 ? 0 : 2 * TILE_BYTES;
  static constexpr int OFF_BAR = OFF_KV + NS * TILE_BYTES;
  static constexpr int NUM_BARS = 4 /*q full,empty*/ + 2 * NS + 2 /*s*/ + 2 /*p*/ + 2 /*o*/ + 4 /*mailbox full,empty*/ + 2 /*q in TMEM (+pad)*/;
  static constexpr int OFF_MBOX = OFF_BAR + NUM_BARS * 8 + 16;      // after the TMEM base slot; 2 x 32 B
  static constexpr int OFF_XCHG = OFF_MBOX + 64;                    // one-tile units: row max / row sum swap area of the warp pairs (2 x 128 floats) + 16 flags
  static constexpr int OFF_STAGE = (OFF_XCHG + (TILES == 1 ? 1024 + 64 : 0) + 1023) / 1024 * 1024;   // epilogue staging: 4 KB per softmax warp, TMA-store sources (1024 B aligned)
  static constexpr int SMEM = OFF_STAGE + 8 * 4096 + 1024;          // +1024: manual alignment
  static_assert(SMEM <= 232448, "shared memory budget");
  static constexpr int TM_S = 0, TM_O = 256;                // TMEM column bases; tile t adds t*128
  // One-tile units (D = 256): the S1 region holds Q as a TMEM A operand (column x = elements 2x, 2x+1 of the
  // row).  QK^T from shared memory needs the full 128 B/clk of shared-memory bandwidth (4 KB of Q + 4 KB of K
  // per 64-clk MMA) and stalls whenever the K/V TMA writes of the next tile land; with Q in TMEM it needs half.
  static constexpr int TM_Q = 128;
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

// What the two Q tiles of a work unit are (d <= 128; above, a unit is one tile):
//  * head pair (GQA with an even group size): the SAME 128 query rows of two heads that share a KV head.  Both tiles
//    then have the same KV range and the same mask, so a causal unit ends for both warpgroups in the same iteration;
//  * row pair (odd group sizes, MHA): 256 consecutive rows of one head.  Under a causal mask the lower tile has one KV
//    tile more than the upper one: that last iteration runs with half the tensor work, and the upper tile's warpgroup
//    waits for it before the issuer starts the next unit (measured: 7000 clk per unit on one warpgroup, 2500 on the other).
// Without a mask the two pairings do the same work and the row pair measured 2 % faster (same-box A/B), so head pairs are
// used under causal / local masks only: 64q/8kv 4 x 2048 causal +8.7 %, 32q/8kv 8 x 2048 causal +2 %, C2 unchanged.
__host__ __device__ __forceinline__ bool tc5_head_pair(const KParams& p) {
#ifdef FA_TC5_NO_HEADPAIR
  return false;
#else
  return p.d <= 128 && (p.group & 1) == 0 && (p.is_causal || p.is_local);
#endif
}
// Query rows per work unit
__host__ __device__ __forceinline__ int tc5_unit_rows(const KParams& p) { return (p.d > 128 || tc5_head_pair(p)) ? TC5_BM : 2 * TC5_BM; }
// Work units per (sequence, m_block) list entry: heads, or head pairs
__host__ __device__ __forceinline__ int tc5_unit_heads(const KParams& p) { return tc5_head_pair(p) ? p.h >> 1 : p.h; }

// One unit of work = 256 (d > 128: 128) query rows of one (sequence, head), or 128 rows of two heads (`head` is the
// first).  The producer warp decodes it and publishes it to the other roles through a 2-deep shared-memory mailbox.
struct Tc5Work { int valid, head, bidb, m_block, sq, sk, q_row0, k_row0; };

__device__ __forceinline__ void tc5_read_work(const uint8_t* mbox, Tc5Work& w) {
  const int4 a = *reinterpret_cast<const int4*>(mbox), b = *reinterpret_cast<const int4*>(mbox + 16);
  w.valid = a.x; w.head = a.y; w.bidb = a.z; w.m_block = a.w; w.sq = b.x; w.sk = b.y; w.q_row0 = b.z; w.k_row0 = b.w;
}

template <typename T, int D, bool HAS_ALIBI, bool PAGED, bool HP>
__global__ void __launch_bounds__(TC5_THREADS, 1)
fa_prefill_tc5_kernel(const __grid_constant__ KParams p, const __grid_constant__ Tc5Maps maps) {
  using Cfg = Tc5Cfg<D>;
  constexpr int ND = Cfg::ND, NS = Cfg::NS;
  constexpr bool IS_BF16 = IsBf16<T>::value;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;   // SWIZZLE_128B needs 1024 B
  uint8_t* smem_al = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t q_sm = smem_base + Cfg::OFF_Q, kv_sm = smem_base + Cfg::OFF_KV;
  const uint32_t bars = smem_base + Cfg::OFF_BAR;
  const uint32_t bar_q = bars;                    // [2] Q_t landed
  const uint32_t bar_q_empty = bars + 16;         // [2] last QK_t of the work unit complete
  const uint32_t bar_kv_full = bars + 32;         // [NS]
  const uint32_t bar_kv_empty = bar_kv_full + 8 * NS;
  const uint32_t bar_s = bar_kv_empty + 8 * NS;   // [2] S_t ready
  const uint32_t bar_p = bar_s + 16;              // [2] P_t written (128 arrivals)
  const uint32_t bar_o = bar_p + 16;              // [2] PV_t(j) complete
  const uint32_t bar_w_full = bar_o + 16;         // [2] mailbox slot published
  const uint32_t bar_w_empty = bar_w_full + 16;   // [2] mailbox slot read by MMA thread + 8 softmax warps
  const uint32_t bar_qt = bar_w_empty + 16;       // [1] one-tile units: Q copied to TMEM by the softmax warps (256 arrivals)
  const uint32_t tmem_slot = bar_qt + 16;
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_al + Cfg::OFF_BAR + Cfg::NUM_BARS * 8);
  uint8_t* mbox = smem_al + Cfg::OFF_MBOX;        // [2][32 B]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  [[maybe_unused]] const bool trace_cta = blockIdx.x == 0 && ((threadIdx.x & 127) == 0 || warp == TC5_MMA_WARP);

  // ---- one-time setup ----
  if (warp == TC5_PRODUCER_WARP && lane == 0) {
    prefetch_tmap(&maps.q); prefetch_tmap(&maps.k); prefetch_tmap(&maps.v); prefetch_tmap(&maps.o);
    for (int t = 0; t < 2; t++) {
      mbar_init(bar_q + 8 * t, 1); mbar_init(bar_q_empty + 8 * t, 1);
      mbar_init(bar_s + 8 * t, 1); mbar_init(bar_p + 8 * t, 128 * (3 - Cfg::TILES)); mbar_init(bar_o + 8 * t, 1);   // one-tile units: both warpgroups publish P_0
      mbar_init(bar_w_full + 8 * t, 1); mbar_init(bar_w_empty + 8 * t, 9);
      mbar_init(bar_qt + 8 * t, 256);
    }
    for (int i = 0; i < NS; i++) { mbar_init(bar_kv_full + 8 * i, 1); mbar_init(bar_kv_empty + 8 * i, 1); }
    if constexpr (Cfg::TILES == 1)      // "check failed" tags of the warp pairs: 0 matches no iteration
      for (int i = 0; i < 16; i++) reinterpret_cast<volatile int*>(smem_al + Cfg::OFF_XCHG + 1024)[i] = 0;
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
    // =============================== scheduler + TMA producer ===============================
    // Work order: p.work_items (built by tc5_schedule_kernel, most expensive first) drawn through a
    // global counter — longest-processing-time-first list scheduling over the persistent CTAs — or,
    // for launches too small to need it, a static decode of (m_block, batch, head).
    constexpr bool hp = HP;                 // head-pair units (tc5_head_pair), chosen by the launcher
    const int unit_rows = hp ? TC5_BM : Cfg::UNIT_ROWS, uh = hp ? p.h >> 1 : p.h;
    const int n_mb = (p.seqlen_q + unit_rows - 1) / unit_rows;
    const int total = p.work_items ? p.work_header[1] * uh : n_mb * p.b * uh;
    const int page = PAGED ? p.page_block_size : TC5_BN;
    const int box_rows = min(page, TC5_BN), nbox = TC5_BN / box_rows;
    int ring = 0;                 // K/V ring cursor, runs on across work units
    int published = 0, fetches = 0;
    // next live work unit: drawn from the sorted list, or decoded arithmetically (static order)
    auto fetch = [&]() {
      Tc5Work wk;
      for (;;) {
        int w = 0;
        if (p.work_items) { if (lane == 0) w = atomicAdd(p.work_header, 1); w = __shfl_sync(0xffffffffu, w, 0); }
        else w = fetches * gridDim.x + ((fetches & 1) ? gridDim.x - 1 - blockIdx.x : blockIdx.x);   // boustrophedon
        fetches++;
        wk.valid = w < total;
        if (!wk.valid) break;
        const int unit = w / uh;
        wk.head = (w - unit * uh) << (hp ? 1 : 0);
        if (p.work_items) { const int2 it = p.work_items[unit]; wk.bidb = it.x; wk.m_block = it.y; }
        else { wk.bidb = unit % p.b; const int mb = unit / p.b; wk.m_block = p.is_causal ? n_mb - 1 - mb : mb; }
        const SeqInfo si(p, wk.bidb);
        wk.sq = si.sq; wk.sk = si.sk;
        wk.q_row0 = p.cu_seqlens_q ? p.cu_seqlens_q[wk.bidb] : 0;
        wk.k_row0 = (!PAGED && p.cu_seqlens_k) ? p.cu_seqlens_k[wk.bidb] : 0;
        const int rows = wk.sq * p.group;
        if (wk.m_block * unit_rows < wk.sq && rows >= p.min_rows && rows <= p.max_rows) break;   // dead units: static order only
      }
      return wk;
    };
    auto publish = [&](const Tc5Work& wk) {
      const int ms = published & 1;
      if (lane == 0) {
        mbar_wait(bar_w_empty + 8 * ms, ((published >> 1) & 1) ^ 1);
        *reinterpret_cast<int4*>(mbox + 32 * ms) = make_int4(wk.valid, wk.head, wk.bidb, wk.m_block);
        *reinterpret_cast<int4*>(mbox + 32 * ms + 16) = make_int4(wk.sq, wk.sk, wk.q_row0, wk.k_row0);
        mbar_arrive(bar_w_full + 8 * ms);
      }
      published++;
    };
    // Unit u+1 is drawn once the loads of unit u are all issued — the ring lets the producer run two KV iterations
    // ahead of the MMAs, which covers the draw (one atomic and three dependent loads), so the other roles still find the
    // next unit in the mailbox at a unit boundary.  Drawing it BEFORE unit u's loads (as this loop first did) commits a
    // CTA to its next unit a whole unit early: with only a few units per CTA (a TP shard: 256 units on 148 CTAs) every
    // CTA then holds two units from the first microsecond on, the cost-sorted list degenerates into a static
    // assignment (measured: 24 KV iterations on the longest CTA, 10 on the shortest, where list scheduling gives 16),
    // and a launch pays two draws in series before its first Q load.
    Tc5Work wk = fetch();
    publish(wk);
    for (int unit_idx = 0; wk.valid; unit_idx++) {
#ifdef FA_TC5_EARLY_DRAW
      const Tc5Work nxt = fetch();
      publish(nxt);
#endif

      const int m0 = wk.m_block * unit_rows, sq = wk.sq, sk = wk.sk;
      const int kvh = wk.head / p.group;
      const TileRange tr0 = tc5_range(p, m0, sq, sk), tr1 = Cfg::TILES > 1 ? (hp ? tr0 : tc5_range(p, m0 + TC5_BM, sq, sk)) : TileRange{0, 0};
      const bool act0 = tr0.je > tr0.jb, act1 = tr1.je > tr1.jb;
      const int j_begin = act0 ? (act1 ? min(tr0.jb, tr1.jb) : tr0.jb) : tr1.jb;
      const int j_end = act0 || act1 ? max(tr0.je, tr1.je) : j_begin;
      if (lane == 0) {
        for (int t = 0; t < 2; t++) {
          if (Cfg::TILES == 1 && t == 0) continue;              // one-tile units: Q_0 is a ring item (below)
          // both Q barriers flip once per work unit (inactive tile: plain arrive), so their parity is the unit's
          mbar_wait(bar_q_empty + 8 * t, (unit_idx & 1) ^ 1);   // the previous unit's last QK_t has read Q_t
          if (!(t == 0 ? act0 : act1)) { mbar_arrive(bar_q + 8 * t); continue; }
          mbar_expect_tx(bar_q + 8 * t, Cfg::TILE_BYTES);
          for (int c = 0; c < ND; c++)
            tma_load_4d(q_sm + t * Cfg::TILE_BYTES + c * Cfg::SUB_BYTES, &maps.q, bar_q + 8 * t,
                        c * 64, wk.head + (hp ? t : 0), wk.q_row0 + m0 + (hp ? 0 : t * TC5_BM), p.cu_seqlens_q ? 0 : wk.bidb);
        }
      }
      __syncwarp();
      if constexpr (Cfg::TILES == 1) {
        if (act0) {     // Q of this unit: next ring slot; warpgroup 1 moves it to TMEM, the issuer releases the slot
          const int slot = ring % NS, round = ring / NS;
          if (lane == 0) {
            mbar_wait(bar_kv_empty + 8 * slot, (round & 1) ^ 1);
            mbar_expect_tx(bar_kv_full + 8 * slot, Cfg::TILE_BYTES);
            for (int c = 0; c < ND; c++)
              tma_load_4d(kv_sm + slot * Cfg::TILE_BYTES + c * Cfg::SUB_BYTES, &maps.q, bar_kv_full + 8 * slot,
                          c * 64, wk.head, wk.q_row0 + m0, p.cu_seqlens_q ? 0 : wk.bidb);
          }
          ring++;
        }
        __syncwarp();
      }
      if constexpr (!PAGED) {
        // Packed varlen K/V (and seqused_k): the V tile that holds the end of the sequence also holds rows of the NEXT
        // sequence (or rows past seqused_k).  P is 0 there, but 0 * NaN/Inf would poison P.V with another request's
        // data, so that one tile is filled by hand — valid rows copied, the rest zeroed, in the swizzle TMA would have
        // used — like the paged tail below (the reference zero-fills them: Clear_OOB_MN, flash_fwd_kernel.h:292).
        // Dense tensors need nothing: the tensor map ends at seqlen_k and TMA zero-fills beyond it.
        // Only the unit's LAST V tile can be such a tile, so the issue loop (one lane) is left as it is and stops one
        // item early when that tile needs the hand fill.
        const int valid_last = sk - (j_end - 1) * TC5_BN;                // rows of the last tile inside the sequence
        const bool tail = j_end > j_begin && valid_last < TC5_BN && (p.cu_seqlens_k != nullptr || p.seqused_k != nullptr);
        if (lane == 0) {
          for (int j = j_begin; j < j_end; j++) {
            for (int kv = 0; kv < 2; kv++) {
              if (tail && kv == 1 && j == j_end - 1) break;
              const int slot = ring % NS, round = ring / NS;
              mbar_wait(bar_kv_empty + 8 * slot, (round & 1) ^ 1);
              mbar_expect_tx(bar_kv_full + 8 * slot, Cfg::TILE_BYTES);
              for (int c = 0; c < ND; c++)
                tma_load_4d(kv_sm + slot * Cfg::TILE_BYTES + c * Cfg::SUB_BYTES, kv == 0 ? &maps.k : &maps.v,
                            bar_kv_full + 8 * slot, c * 64, kvh, wk.k_row0 + j * TC5_BN, p.cu_seqlens_k ? 0 : wk.bidb);
              ring++;
            }
          }
        }
        ring = __shfl_sync(0xffffffffu, ring, 0);
        if (tail) {
          const int slot = ring % NS, round = ring / NS;
          const uint32_t dst = kv_sm + slot * Cfg::TILE_BYTES;
          if (lane == 0) mbar_wait(bar_kv_empty + 8 * slot, (round & 1) ^ 1);
          __syncwarp();
          const T* vrow0 = reinterpret_cast<const T*>(p.v) + (int64_t)kvh * p.v_head_stride +
                           (p.cu_seqlens_k ? (int64_t)0 : (int64_t)wk.bidb * p.v_batch_stride) +
                           (int64_t)(wk.k_row0 + (j_end - 1) * TC5_BN) * p.v_row_stride;
          for (int idx = lane; idx < TC5_BN * ND * 8; idx += 32) {
            const int ch = idx & 7, c = (idx >> 3) % ND, r = (idx >> 3) / ND;
            uint4 val = make_uint4(0, 0, 0, 0);
            if (r < valid_last && c * 64 + ch * 8 < p.d) val = *reinterpret_cast<const uint4*>(vrow0 + (int64_t)r * p.v_row_stride + c * 64 + ch * 8);
            asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dst + c * Cfg::SUB_BYTES + r * 128 + ((ch ^ (r & 7)) << 4)),
                         "r"(val.x), "r"(val.y), "r"(val.z), "r"(val.w) : "memory");
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_kv_full + 8 * slot);           // no TMA bytes outstanding: the plain arrival completes the phase
          ring++;
        }
      } else {
        // Paged K/V (reference src/utils.h:297-315, kernel_traits.h:136-144): a 128-key tile is
        // 128/box pages (or a slice of one big page); lane L looks up the block id of box L in the
        // block table and issues that box's TMA loads itself, so the table indirection costs one
        // coalesced load per tile.  The tensor maps are (d, heads, page_row, block).  The V tile that
        // contains the end of the sequence is special: rows >= seqlen_k hold whatever the pool holds
        // (possibly NaN) and 0 * NaN would poison P.V, so its boxes that reach past the end are copied
        // by hand (valid rows) / zero-filled (the rest) in the 128-B swizzle TMA would have used.
        // single_page without cache_batch_idx: no table, block = sequence index
        const int* table = p.block_table ? p.block_table + (int64_t)wk.bidb * p.block_table_batch_stride : nullptr;
        for (int j = j_begin; j < j_end; j++) {
          const int row0 = j * TC5_BN;
          const int valid = min(TC5_BN, sk - row0);                    // rows of this tile inside the sequence (>= 1)
          for (int kv = 0; kv < 2; kv++, ring++) {
            const int slot = ring % NS, round = ring / NS;
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
                if (r < valid && c * 64 + ch * 8 < p.d) {
                  const int g = row0 + r, pg = g / page;
                  const int blk = table ? table[min(pg, p.max_blocks_per_seq - 1)] : wk.bidb;
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
              const int blk = table ? table[min(pg, p.max_blocks_per_seq - 1)] : wk.bidb;
              for (int c = 0; c < ND; c++)
                tma_load_4d(dst + c * Cfg::SUB_BYTES + lane * box_rows * 128, kv == 0 ? &maps.k : &maps.v, bar,
                            c * 64, kvh, g - pg * page, blk);
            }
          }
        }
      }
#ifndef FA_TC5_EARLY_DRAW
      const Tc5Work nxt = fetch();
      publish(nxt);
#endif
      wk = nxt;
    }
    __syncwarp();
  } else if (warp == TC5_MMA_WARP) {
    setmaxnreg_dec<TC5_REGS_OTHER>();
    // =============================== MMA issuer ===============================
    // One thread issues both tiles' chains in the order PV0(j) QK0(j+1) PV1(j) QK1(j+1): the shared
    // in-order stream is what keeps the two softmax warpgroups in ANTI-phase (one exponentiates
    // while the other tile's MMAs run); two independent issuers drift into phase and halve the
    // MUFU rate each warpgroup sees.  Only one barrier wait (P_t) sits on a tile's chain: the
    // arrival of V_j and K_{j+1} is checked before it.  Work units follow one another without
    // draining: the first QK of the next unit is issued right behind the last PV of this one, so
    // S is ready when a softmax warpgroup returns from its epilogue.
    constexpr uint32_t idesc_qk = make_idesc(IS_BF16, TC5_BM, TC5_BN, false);
    // head dims 32 / 96 run in the D = 64 / 128 instantiation: the TMA boxes clip (zero-fill) at the
    // tensor's true width, QK^T takes d/16 k-steps and PV is issued with N = d, so no tensor-pipe
    // cycle is spent on the padding columns
    const uint32_t idesc_pv = make_idesc(IS_BF16, TC5_BM, p.d, true);
    const int nks = p.d >> 4;
    // descriptor = constant high part | (smem address >> 4); the k-step loops are NOT unrolled: one MMA
    // occupies the tensor pipe for 64 clk, so a few scalar instructions per issue are free, and the
    // issuer's 56 registers hold no precomputed descriptor tables (spills would sit on the P -> PV chain)
    constexpr uint64_t DESC_KMAJOR = ((uint64_t)(16 >> 4) << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
    constexpr uint64_t DESC_V = ((uint64_t)(Cfg::SUB_BYTES >> 4) << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
    if (elect_one()) {
      int ring = 0;               // K/V ring cursor (same sequence of items as the producer's)
      int pbits = 0;              // bit t: parity of the softmax iterations of tile t so far (phase of bar_p)
      // Phase governor.  A tile's chain  softmax -> P seen -> PV + QK -> S seen  is what sets the period, and
      // the two softmax warps of a sub-partition share its MUFU pipe: wherever the two tiles' softmax windows
      // overlap, both slow down by nearly the overlap.  The offset between the tiles is neutrally stable (both
      // pay the same penalty), so nothing pulls it to the overlap-free middle: measured 1320 / 2150 clk instead
      // of 1740 / 1740, softmax 1900 clk instead of 1450.  The issuer therefore holds a tile's MMAs back until
      // half the measured period (minus a margin that lets the estimate decay) has passed since it issued the
      // other tile's.
      [[maybe_unused]] uint32_t gov_t0 = 0, gov_t1 = 0, gov_half = 0;
      [[maybe_unused]] int trace_it = 0;
#ifdef FA_TC5_TRACE
      if (p.trace && trace_cta) {   // SM clock vs wall clock over this CTA's whole life: effective frequency
        long long ns; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns));
        p.trace[(2 * 128 + 127) * 16 + 12] = clock64(); p.trace[(2 * 128 + 127) * 16 + 13] = ns;
      }
      if (p.trace) {                // every CTA: wall-clock start / end, KV iterations, units (behind the CTA-0 timeline)
        long long ns; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns));
        p.trace[3 * 128 * 16 + blockIdx.x * 4 + 0] = ns;
      }
#endif
      for (int wi = 0;; wi++) {
        const int ms = wi & 1;
        mbar_wait(bar_w_full + 8 * ms, (wi >> 1) & 1);
        Tc5Work wk;
        tc5_read_work(mbox + 32 * ms, wk);
        mbar_arrive(bar_w_empty + 8 * ms);
#ifdef FA_TC5_TRACE
        if (!wk.valid && p.trace && trace_cta) {
          long long ns; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns));
          p.trace[(2 * 128 + 127) * 16 + 14] = clock64(); p.trace[(2 * 128 + 127) * 16 + 15] = ns;
          p.trace[(2 * 128 + 127) * 16 + 11] = trace_it;
        }
        if (!wk.valid && p.trace) {
          long long ns; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns));
          p.trace[3 * 128 * 16 + blockIdx.x * 4 + 1] = ns;
          p.trace[3 * 128 * 16 + blockIdx.x * 4 + 2] = trace_it;
          p.trace[3 * 128 * 16 + blockIdx.x * 4 + 3] = wi;
        }
#endif
        if (!wk.valid) break;
        constexpr bool hp = HP;
        const int m0 = wk.m_block * (hp ? TC5_BM : Cfg::UNIT_ROWS);
        const TileRange tr0 = tc5_range(p, m0, wk.sq, wk.sk), tr1 = Cfg::TILES > 1 ? (hp ? tr0 : tc5_range(p, m0 + TC5_BM, wk.sq, wk.sk)) : TileRange{0, 0};
        const bool act0 = tr0.je > tr0.jb, act1 = tr1.je > tr1.jb;
        const int j_begin = act0 ? (act1 ? min(tr0.jb, tr1.jb) : tr0.jb) : tr1.jb;
        const int j_end = act0 || act1 ? max(tr0.je, tr1.je) : j_begin;
        auto active = [&](int t, int j) { return t == 0 ? (j >= tr0.jb && j < tr0.je) : (j >= tr1.jb && j < tr1.je); };
        // S_t = Q_t K_j^T : A = Q (K-major, SW128), B = K (K-major, SW128); D/16 k-steps
        auto issue_qk = [&](int t, int j, int slot, bool release) {
          const uint32_t qa = ((q_sm + t * Cfg::TILE_BYTES) & 0x3FFFF) >> 4, kb = ((kv_sm + slot * Cfg::TILE_BYTES) & 0x3FFFF) >> 4;
#pragma unroll
          for (int ks = 0; ks < D / 16; ks++) {
            if (ks >= nks) break;                       // d = 96 / 32 stop early
            const uint32_t off = (ks >> 2) * (Cfg::SUB_BYTES >> 4) + (ks & 3) * 2;
            if constexpr (Cfg::TILES == 1) umma_ts(tmem + Cfg::TM_S, tmem + Cfg::TM_Q + ks * 8, DESC_KMAJOR | (kb + off), idesc_qk, ks > 0);
            else umma_ss(tmem + Cfg::TM_S + t * 128, DESC_KMAJOR | (qa + off), DESC_KMAJOR | (kb + off), idesc_qk, ks > 0);
          }
          TC5_TRACE(2, trace_it + j - 1 - j_begin, t * 8 + 5);
          umma_commit(bar_s + 8 * t);
          if (release) umma_commit(bar_kv_empty + 8 * slot);
          if (j + 1 == (t == 0 ? tr0.je : tr1.je)) umma_commit(bar_q_empty + 8 * t);   // Q_t may be overwritten
        };
        mbar_wait(Cfg::TILES == 1 ? bar_qt : bar_q, wi & 1);
        mbar_wait(bar_q + 8, wi & 1);
        if constexpr (Cfg::TILES == 1) {
          tc_fence_after();
          if (act0) { mbar_arrive(bar_kv_empty + 8 * (ring % NS)); ring++; }   // Q is in TMEM: its ring slot is free
        }
        if (!act0) mbar_arrive(bar_q_empty);
        if (!act1) mbar_arrive(bar_q_empty + 8);
        if (j_begin < j_end) {
          const int slot = ring % NS;
          mbar_wait(bar_kv_full + 8 * slot, (ring / NS) & 1);            // K_{j_begin}
          tc_fence_after();
          const bool a0 = active(0, j_begin), a1 = active(1, j_begin);
          if (a0) issue_qk(0, j_begin, slot, !a1);
          if (a1) issue_qk(1, j_begin, slot, true);
        }
        for (int j = j_begin; j < j_end; j++) {
          const int vi = ring + 2 * (j - j_begin) + 1, ki = vi + 1;
          const int vslot = vi % NS, kslot = ki % NS;
          const bool more = j + 1 < j_end;
          TC5_TRACE(2, trace_it + j - j_begin, 0);
          mbar_wait(bar_kv_full + 8 * vslot, (vi / NS) & 1);
          if (more) mbar_wait(bar_kv_full + 8 * kslot, (ki / NS) & 1);
          TC5_TRACE(2, trace_it + j - j_begin, 1);
          const bool a0 = active(0, j), a1 = active(1, j);
          const bool k0 = more && active(0, j + 1), k1 = more && active(1, j + 1);
#pragma unroll 1
          for (int t = 0; t < 2; t++) {
            const bool a = t == 0 ? a0 : a1, kq = t == 0 ? k0 : k1;
            const int jb_t = t == 0 ? tr0.jb : tr1.jb;
            if (a) {
              // O_t (+)= P_t V_j : A = P (TMEM), B = V (MN-major, SW128); 8 k-steps of 16 keys
              mbar_wait(bar_p + 8 * t, (uint32_t)(((pbits >> t) ^ (j - jb_t)) & 1));
              TC5_TRACE(2, trace_it + j - j_begin, t * 8 + 2);
#if FA_TC5_GOVERNOR
              if constexpr (Cfg::TILES == 2) {
                if (a0 && a1) {
                  const uint32_t other = t == 0 ? gov_t1 : gov_t0;
                  uint32_t now = clock32();
                  if (gov_half) while ((int32_t)(now - other) < (int32_t)gov_half) now = clock32();
                  if (t == 0) {
                    const uint32_t per = now - gov_t0;              // tile 0 to tile 0
                    if (j > j_begin && per < 16384u) gov_half = (per >> 1) - FA_TC5_GOV_MARGIN;
                    gov_t0 = now;
                  } else gov_t1 = now;
                }
              }
#endif
              tc_fence_after();
              const uint32_t vb = ((kv_sm + vslot * Cfg::TILE_BYTES) & 0x3FFFF) >> 4;
#pragma unroll
              for (int kk = 0; kk < TC5_BN / 16; kk++)
                umma_ts(tmem + Cfg::TM_O + t * 128, tmem + Cfg::TM_S + t * 128 + kk * 8,
                        DESC_V | (vb + kk * (2048 >> 4)), idesc_pv, (j > jb_t) || kk > 0);
              TC5_TRACE(2, trace_it + j - j_begin, t * 8 + 3);
              umma_commit(bar_o + 8 * t);
              if (t == 1 || !a1) umma_commit(bar_kv_empty + 8 * vslot);
              TC5_TRACE(2, trace_it + j - j_begin, t * 8 + 4);
            }
            if (kq) issue_qk(t, j + 1, kslot, t == 1 || !k1);
            TC5_TRACE(2, trace_it + j - j_begin, t * 8 + 6);
          }
          if (!a0 && !a1) umma_commit(bar_kv_empty + 8 * vslot);          // unreachable for monotone windows
          if (more && !k0 && !k1) umma_commit(bar_kv_empty + 8 * kslot);
        }
        ring += 2 * (j_end - j_begin);
        pbits ^= ((tr0.je - tr0.jb) & 1) | (((tr1.je - tr1.jb) & 1) << 1);
#ifdef FA_TC5_TRACE
        trace_it += j_end - j_begin;
#endif
      }
    }
    __syncwarp();
  } else if (warp < 8) {
    setmaxnreg_inc<TC5_REGS_SOFTMAX>();
    // =============================== softmax warpgroups ===============================
    if constexpr (Cfg::TILES == 1) {
#include "fa_prefill_tc5_softmax_pair.inc"
    } else {
#include "fa_prefill_tc5_softmax.inc"
    }
  } else {
    setmaxnreg_dec<TC5_REGS_OTHER>();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == TC5_MMA_WARP) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

// Builds the work list of the persistent kernel: one (sequence, m_block) entry per 256 query rows
// that pass the row filter, ordered by decreasing number of KV tiles (counting sort), and resets
// the draw counter.  header = {counter, n_units}.  One CTA; runs right before the attention kernel.
constexpr int TC5_SCHED_THREADS = 1024, TC5_SCHED_BINS = 2048;
__global__ void __launch_bounds__(TC5_SCHED_THREADS)
tc5_schedule_kernel(const __grid_constant__ KParams p, int2* items, int* header, int max_items) {
  __shared__ int bins[TC5_SCHED_BINS];       // histogram by cost, then running offsets
  __shared__ int warp_sums[32];
  const int tid = threadIdx.x;
  for (int i = tid; i < TC5_SCHED_BINS; i += TC5_SCHED_THREADS) bins[i] = 0;
  __syncthreads();
  const int unit_rows = tc5_unit_rows(p);
  auto cost_of = [&](int mb, int sq, int sk) {
    const TileRange a = tc5_range(p, mb * unit_rows, sq, sk);
    const TileRange b = unit_rows > TC5_BM ? tc5_range(p, mb * unit_rows + TC5_BM, sq, sk) : (tc5_head_pair(p) ? a : TileRange{0, 0});
    return min((a.je - a.jb) + (b.je - b.jb), TC5_SCHED_BINS - 1);
  };
  auto live = [&](int sq) { const int rows = sq * p.group; return sq > 0 && rows >= p.min_rows && rows <= p.max_rows; };
  // The (sequence, m_block) pairs are walked flat — pair idx = s * max_mb + mb with max_mb from the caller's
  // max_seqlen_q — so that a batch of a few long sequences is spread over all threads instead of one thread
  // looping over each sequence's blocks (C2: 8 sequences, up to 32 blocks).  The thread that holds a sequence's
  // last regular block also walks any blocks beyond max_mb (a sequence longer than the stated maximum), and
  // very ragged batches (many sequences, one long) fall back to one thread per sequence.
  const int max_mb = max((p.seqlen_q + unit_rows - 1) / unit_rows, 1);
  const bool flat = (int64_t)p.b * max_mb <= 16 * TC5_SCHED_THREADS;
  const int n_idx = flat ? p.b * max_mb : p.b;
  auto for_each_unit = [&](auto&& fn) {
    for (int idx = tid; idx < n_idx; idx += TC5_SCHED_THREADS) {
      const int s = flat ? idx / max_mb : idx;
      const SeqInfo si(p, s);
      if (!live(si.sq)) continue;
      const int n = (si.sq + unit_rows - 1) / unit_rows;
      if (flat) {
        const int mb = idx - s * max_mb;
        if (mb < n) fn(s, mb, si.sq, si.sk);
        if (mb == max_mb - 1) for (int m2 = max_mb; m2 < n; m2++) fn(s, m2, si.sq, si.sk);
      } else {
        for (int mb = 0; mb < n; mb++) fn(s, mb, si.sq, si.sk);
      }
    }
  };
  for_each_unit([&](int, int mb, int sq, int sk) { atomicAdd(&bins[TC5_SCHED_BINS - 1 - cost_of(mb, sq, sk)], 1); });   // bin 0 = most expensive
  __syncthreads();
  // exclusive prefix sum over the 2048 bins (2 per thread)
  const int a = bins[2 * tid], b = bins[2 * tid + 1];
  int v = a + b;
  const int lane = tid & 31, wp = tid >> 5;
  int incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
  if (lane == 31) warp_sums[wp] = incl;
  __syncthreads();
  if (wp == 0) {
    int s = warp_sums[lane], si = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, si, o); if (lane >= o) si += y; }
    warp_sums[lane] = si - s;
  }
  __syncthreads();
  const int excl = warp_sums[wp] + incl - v;
  bins[2 * tid] = excl; bins[2 * tid + 1] = excl + a;
  if (tid == TC5_SCHED_THREADS - 1) { header[0] = 0; header[1] = min(excl + v, max_items); }
  __syncthreads();
  for_each_unit([&](int s, int mb, int sq, int sk) {
    const int pos = atomicAdd(&bins[TC5_SCHED_BINS - 1 - cost_of(mb, sq, sk)], 1);
    if (pos < max_items) items[pos] = make_int2(s, mb);
  });
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
  // Encoding a map costs ~1-2 us of host time and a prefill launch needs four; a serving loop calls with the same
  // tensors (KV pool, preallocated q / out) over and over, so the last few encodings are kept per host thread, keyed
  // on everything that enters the descriptor.  A CUtensorMap is a plain 128-byte value: a cached copy is as good as
  // a fresh one as long as the key matches.
  struct Key { const void* base; int bf16, box_rows, d, heads, batch; int64_t rows, hs, rs, bs; };
  struct Slot { Key k; CUtensorMap m; bool used; };
  constexpr int NSLOT = 16;
  static thread_local Slot cache[NSLOT] = {};
  static thread_local unsigned next = 0;
  const Key key{base, bf16 ? 1 : 0, box_rows, d, heads, batch, rows, head_stride, row_stride, batch_stride};
  for (int i = 0; i < NSLOT; i++)
    if (cache[i].used && memcmp(&cache[i].k, &key, sizeof(Key)) == 0) { *m = cache[i].m; return true; }
  cuuint64_t dims[4] = {(cuuint64_t)d, (cuuint64_t)heads, (cuuint64_t)rows, (cuuint64_t)batch};
  cuuint64_t strides[3] = {(cuuint64_t)head_stride * 2, (cuuint64_t)row_stride * 2,
                           (cuuint64_t)(batch > 1 ? batch_stride : row_stride * rows) * 2};
  cuuint32_t box[4] = {64, 1, (cuuint32_t)box_rows, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(m, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4,
                  const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return false;
  Slot& s = cache[next++ % NSLOT];
  memset(&s.k, 0, sizeof(Key));          // padding bytes take part in the memcmp
  s.k = key; s.m = *m; s.used = true;
  return true;
}

bool tc5_supported(const KParams& p) {
  if (p.d < 32 || p.d > 256 || p.d % 32 != 0) return false;   // every head dim of the reference build
  if (p.cache_batch_idx) return false;
  if (p.num_splits > 1) return false;
  if (p.single_page) {
    // a non-paged cache seen as one page per sequence (fa_api.cu, as_single_page): no tile straddles a page
    if (p.page_block_size <= 0 || p.num_blocks <= 0 || p.max_blocks_per_seq != 1 || p.cu_seqlens_q) return false;
  } else if (p.block_table) {
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

template <typename T, int D, bool HAS_ALIBI, bool PAGED, bool HP>
static cudaError_t launch_tc5_h(const KParams& p, const Tc5Maps& maps, int grid, cudaStream_t stream) {
  using Cfg = Tc5Cfg<D>;
  auto kern = fa_prefill_tc5_kernel<T, D, HAS_ALIBI, PAGED, HP>;
  FA_SET_SMEM_ONCE(kern, Cfg::SMEM);
  kern<<<grid, TC5_THREADS, Cfg::SMEM, stream>>>(p, maps);
  return cudaGetLastError();
}

template <typename T, int D, bool HAS_ALIBI, bool PAGED>
static cudaError_t launch_tc5_t(const KParams& p, const Tc5Maps& maps, int grid, cudaStream_t stream) {
  if constexpr (D <= 128) { if (tc5_head_pair(p)) return launch_tc5_h<T, D, HAS_ALIBI, PAGED, true>(p, maps, grid, stream); }
  return launch_tc5_h<T, D, HAS_ALIBI, PAGED, false>(p, maps, grid, stream);
}

template <typename T, int D>
static cudaError_t launch_tc5_d(const KParams& p, const Tc5Maps& maps, int grid, cudaStream_t stream) {
  const bool alibi = p.alibi_slopes != nullptr;
  if (p.block_table || p.single_page) return alibi ? launch_tc5_t<T, D, true, true>(p, maps, grid, stream) : launch_tc5_t<T, D, false, true>(p, maps, grid, stream);
  return alibi ? launch_tc5_t<T, D, true, false>(p, maps, grid, stream) : launch_tc5_t<T, D, false, false>(p, maps, grid, stream);
}

// Upper bound on the number of (sequence, m_block) work units, from host-side shapes only.
int64_t tc5_max_units(const KParams& p, int64_t total_q) {
  const int64_t per = tc5_unit_rows(p);
  return p.cu_seqlens_q ? total_q / per + p.b : (int64_t)p.b * ((p.seqlen_q + per - 1) / per);
}
// Workspace for the cost-sorted work list; 0 when the launch is small enough for the static order.
size_t tc5_workspace_bytes(const KParams& p, int64_t total_q, int num_sms) {
  const int64_t units = tc5_max_units(p, total_q);
  if (units * tc5_unit_heads(p) <= num_sms) return 0;
  return 16 + (size_t)units * sizeof(int2);
}

#ifdef FA_TC5_TRACE
// Debug only: a device buffer for the timeline, dumped to $FA_B200_TRACE_FILE after each launch.
static cudaError_t launch_prefill_tc5_impl(const KParams& p, bool is_bf16, int64_t total_q, int64_t total_k, int num_sms, cudaStream_t stream);
cudaError_t launch_prefill_tc5(const KParams& p_in, bool is_bf16, int64_t total_q, int64_t total_k, int num_sms, cudaStream_t stream) {
  static long long* buf = nullptr;
  const size_t n = 3 * 128 * 16 + 4 * 160;
  if (!buf) cudaMalloc(&buf, n * sizeof(long long));
  cudaMemsetAsync(buf, 0, n * sizeof(long long), stream);
  KParams p = p_in; p.trace = buf;
  cudaError_t e = launch_prefill_tc5_impl(p, is_bf16, total_q, total_k, num_sms, stream);
  const char* path = getenv("FA_B200_TRACE_FILE");
  if (e == cudaSuccess && path) {
    std::vector<long long> h(n);
    cudaStreamSynchronize(stream);
    cudaMemcpy(h.data(), buf, n * sizeof(long long), cudaMemcpyDeviceToHost);
    FILE* f = fopen(path, "wb"); if (f) { fwrite(h.data(), sizeof(long long), n, f); fclose(f); }
  }
  return e;
}
static cudaError_t launch_prefill_tc5_impl(const KParams& p, bool is_bf16, int64_t total_q, int64_t total_k, int num_sms, cudaStream_t stream) {
#else
cudaError_t launch_prefill_tc5(const KParams& p, bool is_bf16, int64_t total_q, int64_t total_k, int num_sms, cudaStream_t stream) {
#endif
  Tc5Maps maps;
  const bool varlen = p.cu_seqlens_q != nullptr;
  const int64_t q_rows = varlen ? total_q : p.seqlen_q;
  bool ok = make_map(&maps.q, p.q, is_bf16, TC5_BM, p.d, p.h, q_rows, varlen ? 1 : p.b, p.q_head_stride, p.q_row_stride, p.q_batch_stride) &&
            make_map(&maps.o, p.o, is_bf16, 32, p.d, p.h, q_rows, varlen ? 1 : p.b, p.o_head_stride, p.o_row_stride, p.o_batch_stride);
  if (p.block_table || p.single_page) {   // (d, heads, row in page, block)
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
  const int64_t units = tc5_max_units(p, total_q);
  int grid = (int)std::min<int64_t>(units * tc5_unit_heads(p), num_sms);
  if (grid <= 0) return cudaSuccess;
  if (p.work_items) {
    tc5_schedule_kernel<<<1, TC5_SCHED_THREADS, 0, stream>>>(p, p.work_items, p.work_header, (int)units);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    grid = num_sms;
  }
  if (is_bf16) return p.d <= 64 ? launch_tc5_d<__nv_bfloat16, 64>(p, maps, grid, stream)
                    : p.d <= 128 ? launch_tc5_d<__nv_bfloat16, 128>(p, maps, grid, stream) : launch_tc5_d<__nv_bfloat16, 256>(p, maps, grid, stream);
  return p.d <= 64 ? launch_tc5_d<__half, 64>(p, maps, grid, stream)
       : p.d <= 128 ? launch_tc5_d<__half, 128>(p, maps, grid, stream) : launch_tc5_d<__half, 256>(p, maps, grid, stream);
}

}  // namespace fa
