// fa_api.cu — the C ABI of libfa_b200.so (include/fa_b200.h): argument validation, the
// reference's parameter normalisation, kernel choice, split-KV heuristic and launches.
// This file is the B200-native replacement of the reference's host glue
// (csrc/flash_attn/flash_api.cpp:21-937) minus everything that needs torch (allocation and
// tensor unwrapping live in the Python shim, vllm_flash_attn/_cabi.py).
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <climits>
#include <cstdlib>
#include <cstring>
#include <map>
#include <utility>
#include <vector>
#include "fa_b200.h"
#include "fa_launch.h"

using namespace fa;

static thread_local char g_err[768] = "";
static int fail(int code, const char* fmt, ...) {
  va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof g_err, fmt, ap); va_end(ap);
  return code;
}
#define FA_CHECK(cond, ...) do { if (!(cond)) return fail(FA_ERR_INVALID, __VA_ARGS__); } while (0)

extern "C" int fa_b200_abi_version(void) { return FA_B200_ABI_VERSION; }
extern "C" size_t fa_b200_params_size(void) { return sizeof(FaFwdParams); }
extern "C" const char* fa_b200_last_error(void) { return g_err; }

// Wave-efficiency split choice — the rule of num_splits_heuristic (flash_api.cpp:158-192):
// 1 split if the grid nearly fills the SMs; otherwise the smallest eligible split count whose
// wave efficiency is within 85 % of the best.
extern "C" int fa_b200_num_splits_heuristic(int work, int num_sms, int num_n_blocks, int max_splits) {
  if (work >= 0.8f * num_sms) return 1;
  max_splits = std::min(std::min(max_splits, num_sms), num_n_blocks);
  if (max_splits < 1) return 1;
  auto cdiv = [](int a, int b) { return (a + b - 1) / b; };
  auto eligible = [&](int s) { return s == 1 || cdiv(num_n_blocks, s) != cdiv(num_n_blocks, s - 1); };
  std::vector<float> eff(max_splits + 1, 0.f);
  float best = 0.f;
  for (int s = 1; s <= max_splits; s++) {
    if (!eligible(s)) continue;
    float waves = float(work * s) / num_sms;
    eff[s] = waves / std::ceil(waves);
    best = std::max(best, eff[s]);
  }
  for (int s = 1; s <= max_splits; s++)
    if (eligible(s) && eff[s] >= 0.85f * best) return s;
  return 1;
}

// Decode kernel, every shape that decode_choose() below does not cover (d != 128, 17-32 packed rows, sharded calls): the
// rule above with two corrections measured on B200 (profiles/r02_decode_split_sweep.jsonl: explicit
// split counts against the automatic one, 20-256 sequences x 8 kv heads; slots = 2 resident CTAs per SM, blocks of 256
// keys).  (1) 1.0-1.5 waves of units (the reference never splits above 0.8 waves): the second wave runs almost empty, so
// 2-4 pieces per unit are taken when they fill the waves better AND a piece keeps >= 1024 keys (48 sequences: 4096 keys
// 168 -> 149 us, 16384 keys 630 -> 516 us).  (2) Pieces shorter than 1024 keys must not spill into a second wave — a
// piece's fixed cost (prologue chain, partial write, ticket, merge) is then as long as its stream (24 sequences x 1024
// keys: 4 splits 38.5 us, 1 split 26.7 us; 20 sequences x 4096 keys: 8 splits -> 3, 73.9 us).
static int decode_splits(int work, int slots, int n_blocks) {
  constexpr int kMinBlocks = 4;
  auto eff = [&](int s) { const float w = float(work) * s / slots; return w / std::ceil(w); };
  int s = fa_b200_num_splits_heuristic(work, slots, n_blocks, 128);
  if (s == 1 && work > slots && eff(1) < 0.75f) {
    float best = eff(1);
    for (int t = 2; t <= 4; t++) if (n_blocks / t >= kMinBlocks) best = std::max(best, eff(t));
    for (int t = 1; t <= 4; t++) if ((t == 1 || n_blocks / t >= kMinBlocks) && eff(t) >= 0.85f * best) { s = t; break; }
  }
  auto fits = [&](int t) { return t == 1 || n_blocks / t >= kMinBlocks || (int64_t)work * t <= slots; };
  if (!fits(s)) {                              // the best-filling count below it that keeps its pieces long or in one wave
    int bs = 1; float be = eff(1);
    for (int t = 2; t < s; t++) if (fits(t) && eff(t) > be + 1e-3f) { be = eff(t); bs = t; }
    s = bs;
  }
  return s;
}

// d = 128, up to 16 packed rows: the decode kernel has two grid forms — 4 warps x 3 stages at 2 CTAs per SM and 4 x 2 at
// 3 CTAs per SM (fa_decode.cu) — and which is faster depends on how units x splits falls on 2 or 3 slots per SM, so form
// and split count are chosen TOGETHER by a small time model instead of the wave rule alone.  A launch is waves of
// resident CTAs; in a wave of n CTAs a piece streams at min(what one CTA reaches with c CTAs on its SM, HBM / n); every
// wave pays a fixed cost, a split launch a merge.  Constants fitted to a sweep of both forms x 7 split counts x 42
// shapes on B200 (32-1024 units, 1024-16384 keys; profiles/r02_decode_form_sweep.jsonl): the model's pick is the
// measured best or within 4.1 % of it (mean 0.5 %); against the wave rule it gains 8-25 % at 160-400 and 640-770
// units (40 sequences x 8 kv heads x 1024 keys: 52.1 -> 39.3 us) and chooses the same counts for small batches.
struct DecodeChoice { int form, splits; };
static DecodeChoice decode_choose(int64_t units, int seqlen_k, int num_sms, bool allow_three_per_sm) {
  static const float kCtaRate[2][3] = {{34.f, 27.f, 27.f}, {34.f, 27.f, 20.f}};   // GB/s per CTA with 1 / 2 / 3 CTAs on the SM
  const float kHbm = 6900.f, kFix[2] = {3.0f, 3.5f}, kBase = 6.f, kMerge = 2.f, kPerSplit = 0.15f;   // GB/s, us
  const int n_blocks = std::max(1, (seqlen_k + 255) / 256);
  const float unit_bytes = (float)seqlen_k * 512.f;                                // K + V rows of 256 B
  DecodeChoice best{0, 1}; float best_t = 1e30f;
  for (int s = 1; s <= std::min(128, n_blocks); s++)
    for (int f = 0; f < (allow_three_per_sm ? 2 : 1); f++) {
      const int64_t slots = (int64_t)(f ? 3 : 2) * num_sms, pieces = units * s;
      const float piece_bytes = unit_bytes / s;
      auto wave = [&](int64_t n) {
        const int c = (int)std::min<int64_t>((n + num_sms - 1) / num_sms, f ? 3 : 2);
        const float rate = std::min(kCtaRate[f][c - 1], kHbm / (float)n);
        return piece_bytes / rate * 1e-3f + kFix[f];
      };
      float t = kBase + (s > 1 ? kMerge + kPerSplit * s : 0.f);
      if (pieces / slots) t += (float)(pieces / slots) * wave(slots);
      if (pieces % slots) t += wave(pieces % slots);
      if (t < best_t - 1e-4f) { best_t = t; best = DecodeChoice{f, s}; }
    }
  return best;
}

static int device_sms() {
  static int sms[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev < 0 || dev >= 64) dev = 0;
  if (!sms[dev]) { int v = 0; cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev); sms[dev] = v > 0 ? v : 148; }
  return sms[dev];
}

enum Op { OP_FWD = 0, OP_VARLEN = 1, OP_KVCACHE = 2 };

// fwd_kvcache over a NON-paged cache with per-sequence lengths and / or cache_batch_idx (flash_api.cpp:859-901) on the
// tcgen05 prefill kernel: the cache is a page pool whose pages are whole cache entries (seqlen_k rows, one page per
// sequence, so no tile straddles a page) and cache_batch_idx is its one-column block table; the paged path's V-tail fill
// keeps rows past cache_seqlens (which may hold anything) out of P.V.  Rewrites *k on success.
static bool as_single_page(const FaFwdParams* a, int op, KParams* k) {
  if (op != OP_KVCACHE || a->block_table || a->num_blocks <= 0 || a->seqlen_k <= 0) return false;
  if (!a->cache_batch_idx && !(a->cu_seqlens_k && !k->is_seqlens_k_cumulative)) return false;
  KParams t = *k;
  t.single_page = 1;
  t.block_table = a->cache_batch_idx; t.block_table_batch_stride = 1;
  t.page_block_size = a->seqlen_k; t.max_blocks_per_seq = 1; t.num_blocks = a->num_blocks;
  t.cache_batch_idx = nullptr;
  if (!tc5_supported(t)) return false;
  *k = t;
  return true;
}

// Side stream + fork / join events for the decode half of a mixed varlen batch, one set per (host thread, device,
// caller stream).  Created on first use and kept for the life of the process (a handful of streams in practice).
// The fork / join pattern is capture-safe: under stream capture the side stream joins the capture at the fork
// event and leaves it at the join event.
// Balanced-range decode (fa_decode_streamk_kernel).  Measured on B200 against the classic grid with the heuristic's split
// count (profiles/r02_decode_streamk.txt): it wins where one CTA per (unit, equal piece) quantises badly against the 296
// CTA slots or the lengths are ragged — the decode half of C4's TP shards: 192 sequences of 1024..4096 keys, 65.0 -> 58.5 us
// at 1 KV head, 108 -> 98 us at 2 — and loses 5..10 us where the classic grid already fits (a CTA whose range crosses a
// unit boundary pays the per-piece prologue and merge twice).  Inside the whole mixed call the gain does not survive
// either: C4's TP=8 shard 0.1269 -> 0.1233 ms eager but 0.1178 -> 0.1272 ms under a CUDA graph — the classic grid's many
// small CTAs drain unevenly and the prefill CTAs (which cannot share an SM with a decode CTA) start in the gaps, the
// balanced grid ends all at once.  So the form is OPT-IN: FA_B200_STREAMK=1 in the environment turns it on wherever it
// is applicable (ragged decode batches, very large totals, development); default off.
static int streamk_mode() {
  static const int mode = [] { const char* e = getenv("FA_B200_STREAMK"); return e ? (e[0] == '0' ? 0 : 1) : -1; }();
  return mode;
}
static bool streamk_wanted(int64_t units, int num_sms, bool mixed_half) {
  const int m = streamk_mode();
  (void)units; (void)num_sms; (void)mixed_half;
  return m == 1;
}
struct ForkSet { cudaStream_t side, side2; cudaEvent_t fork, join, join2; };
static ForkSet* get_fork_set(cudaStream_t main) {
  // per host thread: two threads that issue on the same stream must not share the fork / join events (an event records
  // the LAST cudaEventRecord: thread B's record between thread A's record and wait would let A's side stream start early)
  static thread_local std::map<std::pair<int, cudaStream_t>, ForkSet> sets;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return nullptr;
  auto key = std::make_pair(dev, main);
  auto it = sets.find(key);
  if (it != sets.end()) return &it->second;
  ForkSet f;
  if (cudaStreamCreateWithFlags(&f.side, cudaStreamNonBlocking) != cudaSuccess ||
      cudaStreamCreateWithFlags(&f.side2, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
  if (cudaEventCreateWithFlags(&f.fork, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&f.join, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&f.join2, cudaEventDisableTiming) != cudaSuccess) return nullptr;
  return &sets.emplace(key, f).first->second;
}

struct Plan {
  KParams k;
  int kernel;      // FA_KERNEL_*
  int num_splits;
  size_t ws_bytes;
  bool is_bf16;
  int num_sms;
  int decode_rows;   // > 0: mixed varlen batch — sequences with at most this many packed rows go to the decode kernel
  bool fused_tail;   // decode kernel merges its own splits (and exchanges across GPUs): no separate combine launch
  size_t lse_floats, o_floats;   // split workspace payload: lse block (padded to 4 floats) | o block | [arrival counters]
  bool counters_in_head;
  bool streamk;                // decode kernel in its balanced-range form (fa_decode_streamk_kernel)
  int sk_slots;                // ... piece slots of its partial pool
  int mixed_splits;            // mixed varlen batch: split count of the decode half (>= 1)
  int mixed_form;              // ... and its grid form (KParams::wide_cta of the 16-row launch: 0 or 2)
  size_t mixed_part_off;       // ... byte offset of its partials in the workspace (behind the prefill work list)
  size_t mixed_lse_floats;
};
static_assert(FA_MAX_PEERS == FA_KP_MAX_PEERS, "peer capacity of the C ABI and of KParams must agree");

// Validation + normalisation shared by the three ops.  Mirrors the checks of mha_fwd /
// mha_varlen_fwd / mha_fwd_kvcache that do not need tensor metadata beyond what FaFwdParams holds.
static int make_plan(const FaFwdParams* a, int op, Plan* pl) {
  FA_CHECK(a != nullptr, "params is NULL");
  FA_CHECK(a->q && a->k && a->v && a->o, "q, k, v, out must be non-NULL device pointers");
  FA_CHECK(a->b > 0, "batch size must be positive");                                   // flash_api.cpp:284
  FA_CHECK(a->d > 0 && a->d <= 256, "FlashAttention forward only supports head dimension at most 256");  // :286
  FA_CHECK(a->d % 8 == 0, "head dimension must be a multiple of 8 (pad in the caller, flash_api.cpp:310-319)");
  FA_CHECK(a->h > 0 && a->h_k > 0 && a->h % a->h_k == 0,
           "Number of heads in key/value must divide number of heads in query");       // :287
  FA_CHECK(a->seqlen_q >= 0 && a->seqlen_k >= 0, "negative sequence length");
  FA_CHECK(a->is_bf16 == 0 || a->is_bf16 == 1, "FlashAttention only support fp16 and bf16 data type");  // :262
  if (op == OP_VARLEN) FA_CHECK(a->cu_seqlens_q && a->cu_seqlens_k, "cu_seqlens_q/k must be given for varlen_fwd");
  if (op != OP_VARLEN) FA_CHECK(a->cu_seqlens_q == nullptr, "cu_seqlens_q is only valid for varlen_fwd");
  const bool paged = a->block_table != nullptr;
  if (paged) {
    FA_CHECK(op != OP_FWD, "block_table is not supported by fwd");
    FA_CHECK(a->page_block_size > 0 && a->page_block_size % 16 == 0,
             "Paged KV cache block size must be divisible by 16");                     // :491,:735
    FA_CHECK(a->cache_batch_idx == nullptr, "Paged KVcache does not support cache_batch_idx");  // :718
    FA_CHECK(a->max_blocks_per_seq > 0, "block_table must have at least one column");
  }
  if (a->alibi_slopes == nullptr) FA_CHECK(a->alibi_slopes_batch_stride == 0, "alibi stride without slopes");
  // New K/V rows given to the attention call itself: appended (and rotated, with q) INSIDE the decode kernel — the
  // reference's fused path (flash_fwd_kernel.h:672-824).  Only fwd_kvcache, only when the call runs on the decode
  // kernel (checked below, fa_b200_fused_append_supported tells beforehand); otherwise the caller runs
  // fa_b200_append_kv first and passes seqlen_knew alone.
  const bool fused_append = a->knew != nullptr;
  if (fused_append) {
    FA_CHECK(op == OP_KVCACHE, "knew / vnew are only valid for fwd_kvcache");
    FA_CHECK(a->vnew != nullptr, "key and value must be passed together");
    FA_CHECK(a->cu_seqlens_k != nullptr, "seqlens_k must be passed if key/value are passed");
    FA_CHECK(a->seqlen_knew > 0, "seqlen_knew must be positive when knew is given");
    if (a->rotary_cos) {
      FA_CHECK(a->rotary_sin != nullptr, "If rotary cos is provided, rotary sin must also be provided");
      FA_CHECK(a->rotary_dim <= a->d, "rotary_dim must be <= headdim");
      FA_CHECK(a->rotary_dim % 16 == 0, "Only rotary dimensions divisible by 16 are currently supported");
    }
  } else {
    FA_CHECK(a->vnew == nullptr && a->rotary_cos == nullptr,
             "rotary / vnew without knew: append-KV and rotary are applied by fa_b200_append_kv or by passing knew to fwd_kvcache");
  }
  FA_CHECK(a->seqlen_knew >= 0 && (a->seqlen_knew == 0 || op == OP_KVCACHE), "seqlen_knew is only valid for fwd_kvcache");

  KParams& k = pl->k;
  memset(&k, 0, sizeof k);
  k.q = a->q; k.k = a->k; k.v = a->v; k.o = a->o; k.lse = a->softmax_lse;
  k.q_batch_stride = a->q_batch_stride; k.q_row_stride = a->q_row_stride; k.q_head_stride = a->q_head_stride;
  k.k_batch_stride = a->k_batch_stride; k.k_row_stride = a->k_row_stride; k.k_head_stride = a->k_head_stride;
  k.v_batch_stride = a->v_batch_stride; k.v_row_stride = a->v_row_stride; k.v_head_stride = a->v_head_stride;
  k.o_batch_stride = a->o_batch_stride; k.o_row_stride = a->o_row_stride; k.o_head_stride = a->o_head_stride;
  k.b = a->b; k.seqlen_q = a->seqlen_q; k.seqlen_k = a->seqlen_k; k.h = a->h; k.h_k = a->h_k; k.d = a->d;
  k.group = a->h / a->h_k;
  k.scale = a->softmax_scale; k.scale_log2 = a->softmax_scale * 1.4426950408889634f;
  k.alibi_slopes = a->alibi_slopes; k.alibi_batch_stride = a->alibi_slopes_batch_stride;
  k.cu_seqlens_q = a->cu_seqlens_q; k.cu_seqlens_k = a->cu_seqlens_k; k.seqused_k = a->seqused_k;
  k.is_seqlens_k_cumulative = op == OP_KVCACHE ? 0 : 1;
  if (op == OP_KVCACHE && a->cu_seqlens_k == nullptr) k.is_seqlens_k_cumulative = 1;
  k.block_table = a->block_table; k.block_table_batch_stride = a->block_table_batch_stride;
  k.page_block_size = a->page_block_size; k.max_blocks_per_seq = a->max_blocks_per_seq; k.num_blocks = a->num_blocks;
  k.cache_batch_idx = a->cache_batch_idx;
  k.seqlen_knew = a->seqlen_knew;   // rows already appended by fa_b200_append_kv
  k.lse_seq_stride = a->seqlen_q;
  k.part_seq_stride = a->seqlen_q;
  k.min_rows = 0; k.max_rows = INT_MAX;
  FA_CHECK(a->kv_shard_len >= 0 && a->kv_shard_start >= 0 && (a->kv_shard_len == 0 || op == OP_KVCACHE), "kv_shard_* is only valid for fwd_kvcache");
  k.kv_shard_start = a->kv_shard_start; k.kv_shard_len = a->kv_shard_len;
  k.partial = 0;
  if (a->partial_out) {
    FA_CHECK(op == OP_KVCACHE, "partial_out is only valid for fwd_kvcache");
    FA_CHECK(a->softmax_lse != nullptr, "partial_out needs softmax_lse");
    FA_CHECK(a->o_row_stride == a->d && a->o_head_stride == (int64_t)a->seqlen_q * a->d &&
             a->o_batch_stride == (int64_t)a->h * a->seqlen_q * a->d,
             "partial_out: o must be fp32 (b,h,seqlen_q,d) contiguous");
    FA_CHECK((reinterpret_cast<uintptr_t>(a->o) & 15) == 0, "partial_out: o must be 16-byte aligned");
    k.partial = 1;
  }

  // ---- mask normalisation (flash_api.cpp:289-294 and set_params_fprop :119-126) ----
  int wl = a->window_left, wr = a->window_right;
  bool causal = a->is_causal != 0;
  if (wl >= a->seqlen_k) wl = -1;
  if (wr >= a->seqlen_k) wr = -1;
  if (a->seqlen_q == 1 && !a->alibi_slopes) causal = false;   // :293,:493,:744
  if (causal) wr = 0;
  k.is_causal = wl < 0 && wr == 0;
  if (wl < 0 && wr >= 0) wl = a->seqlen_k;
  if (wl >= 0 && wr < 0) wr = a->seqlen_k;
  k.window_left = wl; k.window_right = wr;
  k.is_local = (wl >= 0 || wr >= 0) && !k.is_causal;

  // A KV shard knows its keys only by LOCAL position: masks and biases that depend on the global key index would be
  // silently wrong, and without a partial result there is nothing to merge with the other shards.
  if (a->kv_shard_len > 0) {
    FA_CHECK(a->alibi_slopes == nullptr, "kv_shard: ALiBi is not supported (the bias depends on the global key position)");
    FA_CHECK(a->seqlen_q == 1 || !(k.is_causal || k.is_local), "kv_shard: causal / local masks need seqlen_q == 1");
    FA_CHECK(a->partial_out || a->fused_exchange, "kv_shard: needs partial_out or fused_exchange (a shard's result is a partial)");
  }
  const FaFusedExchange* fx = a->fused_exchange;
  if (fx) {
    FA_CHECK(op == OP_KVCACHE && a->kv_shard_len > 0 && !a->partial_out, "fused_exchange: fwd_kvcache with kv_shard_len > 0 and without partial_out only");
    FA_CHECK(fx->world >= 1 && fx->world <= FA_MAX_PEERS && fx->rank >= 0 && fx->rank < fx->world, "fused_exchange: bad rank/world");
    FA_CHECK(fx->state && fx->split_counters, "fused_exchange: NULL state / split_counters");
    for (int r = 0; r < fx->world; r++)
      FA_CHECK(fx->recv[r] && fx->flags[r] && (reinterpret_cast<uintptr_t>(fx->recv[r]) & 15) == 0, "fused_exchange: recv/flags pointer missing or misaligned");
    FA_CHECK((int64_t)a->b * a->h_k <= fx->max_units, "fused_exchange: b*h_k = %lld exceeds max_units = %d", (long long)a->b * a->h_k, fx->max_units);
    FA_CHECK(fx->unit_elems % 4 == 0 && fx->unit_elems >= (int64_t)a->seqlen_q * k.group * (a->d + 1),
             "fused_exchange: unit_elems too small for seqlen_q*group*(d+1) or not a multiple of 4");
    FA_CHECK(a->softmax_lse != nullptr, "fused_exchange needs softmax_lse");
  }

  pl->is_bf16 = a->is_bf16 != 0;
  const int num_sms = a->num_sms > 0 ? a->num_sms : device_sms();

  // ---- kernel choice ----
  const int rows = a->seqlen_q * k.group;     // packed rows per (sequence, kv head)
  int kernel = a->force_kernel;
  k.num_splits = 1;
  pl->decode_rows = 0;
  if (kernel == FA_KERNEL_NONE) {
    if (decode_supported(a->d, rows)) kernel = FA_KERNEL_DECODE;
    else if (!k.partial && (tc5_supported(k) || as_single_page(a, op, &k))) {
      kernel = FA_KERNEL_TC5_PREFILL;
      // A varlen batch may mix prompts with single-token (or few-token) sequences — vLLM's chunked
      // prefill.  Those would waste a 256-row tile per head, so they are routed to the decode
      // kernel (one CTA per KV head, K/V read once per group) by a second launch over the same batch.
      if (op == OP_VARLEN && decode_supported(a->d, k.group)) pl->decode_rows = a->d <= 128 ? 32 : 16;   // the decode kernel's row capacity
    } else kernel = FA_KERNEL_TILE;
  }
  if (fused_append) {
    FA_CHECK(kernel == FA_KERNEL_DECODE && !k.is_local && a->kv_shard_len == 0,
             "in-kernel append needs a decode shape (seqlen_q*(h/h_k) <= 32; 16 above d = 128), no sliding window and no kv_shard: "
             "run fa_b200_append_kv first and pass seqlen_knew alone (fa_b200_fused_append_supported tells beforehand)");
    k.knew = a->knew; k.vnew = a->vnew;
    k.knew_batch_stride = a->knew_batch_stride; k.knew_row_stride = a->knew_row_stride; k.knew_head_stride = a->knew_head_stride;
    k.vnew_batch_stride = a->vnew_batch_stride; k.vnew_row_stride = a->vnew_row_stride; k.vnew_head_stride = a->vnew_head_stride;
    k.rotary_cos = a->rotary_cos; k.rotary_sin = a->rotary_sin;
    k.rotary_dim = a->rotary_cos ? a->rotary_dim : 0; k.seqlen_ro = a->seqlen_ro; k.rotary_interleaved = a->is_rotary_interleaved;
    k.rot_by_qpos = (k.is_causal || k.is_local) ? 1 : 0;       // q row i rotated at cache_len + i (flash_fwd_kernel.h:794)
  }
  if (fx) FA_CHECK(kernel == FA_KERNEL_DECODE, "fused_exchange: the call must have a decode shape (seqlen_q*(h/h_k) <= 32; 16 above d = 128)");
  if (kernel == FA_KERNEL_DECODE) FA_CHECK(decode_supported(a->d, rows), "decode kernel: d/rows not supported");
  if (kernel == FA_KERNEL_TC5_PREFILL) FA_CHECK(!k.partial && (tc5_supported(k) || as_single_page(a, op, &k)), "tcgen05 prefill kernel: shape/layout not supported");
  if (kernel == FA_KERNEL_TILE) FA_CHECK(tile_supported(a->d), "head dimension must be a multiple of 32 (reference build: -DFLASHATTENTION_DISABLE_UNEVEN_K)");
  pl->kernel = kernel;

  // ---- split-KV choice ----
  // The reference consults its heuristic in mha_fwd (:375-377), in varlen only for GQA-packed
  // decode (:618-623) and in kvcache with the caller's num_splits (:903-905).  We split only in
  // the two kernels that write partials, sized for 2 resident CTAs per SM.
  int splits = 1, decode_form = 0;
  const bool may_split = kernel != FA_KERNEL_TC5_PREFILL &&
                         (op == OP_KVCACHE || op == OP_FWD || (op == OP_VARLEN && a->seqlen_q == 1));
  if (may_split && a->seqlen_k > 0) {
    const int req = op == OP_KVCACHE ? a->num_splits : 0;
    if (req >= 1) splits = req;
    else {
      const int blk = kernel == FA_KERNEL_DECODE ? 256 : 128;   // keys per unit of split work
      const int n_blocks = (a->seqlen_k + blk - 1) / blk;
      const int work = kernel == FA_KERNEL_DECODE ? a->b * a->h_k
                                                  : a->b * a->h * ((a->seqlen_q + 63) / 64);
      if (kernel == FA_KERNEL_DECODE && a->d == 128 && rows <= 16 && !fx) {
        const DecodeChoice c = decode_choose(work, a->seqlen_k, num_sms, true);
        splits = c.splits; decode_form = c.form;
      } else
        splits = kernel == FA_KERNEL_DECODE ? decode_splits(work, num_sms * 2, n_blocks)
                                            : fa_b200_num_splits_heuristic(work, num_sms * 2, n_blocks, 128);
    }
    FA_CHECK(splits <= 128, "num_splits > 128 not supported");                         // :217
  }
  // Few units (small decode batches, TP shards with one or two KV heads): the decode kernel's balanced-range form — a
  // fixed grid of 2 CTAs per SM, every CTA an equal share of ALL units' chunks (fa_decode.cu).  Taken whenever the
  // split count is ours to choose; a caller's explicit num_splits keeps the classic grid.
  pl->streamk = false; pl->sk_slots = 0;
#ifndef FA_NO_STREAMK
  if (kernel == FA_KERNEL_DECODE && may_split && a->seqlen_k > 0 && !(op == OP_KVCACHE && a->num_splits >= 1) &&
      a->b <= 512 && (int64_t)a->b * a->h_k < 4 * (int64_t)num_sms * 2 && (size_t)a->b * a->h_k * sizeof(int) <= FA_B200_WS_HEAD_BYTES &&
      streamk_wanted((int64_t)a->b * a->h_k, num_sms, false)) {
    pl->streamk = true;
    k.sk_grid = num_sms * 2; k.sk_min_per_cta = 8;
    k.sk_rows = rows <= 16 ? 16 : 32;
    pl->sk_slots = k.sk_grid + a->b * a->h_k;
    const int max_chunks = (a->seqlen_k + 15) / 16;
    splits = std::max(2, std::min(128, std::min(k.sk_grid, max_chunks)));     // reported as num_splits: pieces a unit can have at most
  }
#endif
  k.num_splits = pl->num_splits = splits;
  pl->num_sms = num_sms;
  // one wave of long pieces: the 12-warp form of the d = 128 decode kernel (fa_decode.cu launch_decode_d).  seqlen_k is the
  // cache's capacity (block-table width), not the lengths on the device: a wide table over short contexts picks it too
  // and pays its larger fixed cost (8 sequences x 1024 keys: 16.4 -> 18.2 us), hence the high bar of 4096 keys per piece.
  k.wide_cta = kernel == FA_KERNEL_DECODE && a->d == 128 && rows <= 16 && !pl->streamk && splits >= 1 &&
               (int64_t)a->b * a->h_k * splits <= 2 * (int64_t)num_sms && a->seqlen_k / splits >= 4096;
  if (decode_form == 1 && !pl->streamk) k.wide_cta = 2;                     // three smaller CTAs per SM (decode_choose)
  // split workspace: lse block, padded to 4 floats so that the o block stays 16-byte aligned | o block | the arrival
  // counters of the decode kernel's in-kernel merge (when they are not part of a fused_exchange allocation)
  const size_t part_rows = pl->streamk ? (size_t)pl->sk_slots * k.sk_rows : (size_t)splits * a->b * a->h * std::max(a->seqlen_q, 1);
  pl->lse_floats = (part_rows + 3) / 4 * 4;
  pl->o_floats = part_rows * a->d;
  pl->fused_tail = kernel == FA_KERNEL_DECODE && (splits > 1 || fx != nullptr);
  // arrival counters: in the workspace head (kept at zero between calls: workspace_clean) when they fit, else behind the partials
  pl->counters_in_head = (size_t)a->b * a->h_k * sizeof(int) <= FA_B200_WS_HEAD_BYTES;
  pl->ws_bytes = splits > 1 ? FA_B200_WS_HEAD_BYTES + (pl->lse_floats + pl->o_floats) * sizeof(float) +
                              ((pl->fused_tail && !fx && !pl->counters_in_head) ? (size_t)a->b * a->h_k * sizeof(int) : 0) : 0;
  pl->mixed_splits = 1; pl->mixed_part_off = 0; pl->mixed_lse_floats = 0; pl->mixed_form = 0;
  if (kernel == FA_KERNEL_TC5_PREFILL) {
    if (op == OP_VARLEN) FA_CHECK(a->total_q > 0, "total_q (rows of q) must be given for varlen_fwd");
    pl->ws_bytes = tc5_workspace_bytes(k, a->total_q, num_sms);    // cost-sorted work list of the persistent kernel
    if (pl->ws_bytes) pl->ws_bytes += FA_B200_WS_HEAD_BYTES;
    // Decode half of a mixed batch: with few (sequence, kv head) units — a TP shard holding one or two KV heads — one CTA
    // per unit leaves SMs idle and the longest context sets the time, so the half is split like a kvcache decode
    // (the reference consults the same heuristic for its GQA-packed varlen decode, flash_api.cpp:618-623).  Which
    // sequences are short is only known on the device: the count is taken as if all were.
    pl->sk_slots = 0;
    if (pl->decode_rows > 0 && a->seqlen_k > 0 && (size_t)a->b * a->h_k * sizeof(int) <= FA_B200_WS_HEAD_BYTES) {
      int ms = fa_b200_num_splits_heuristic(a->b * a->h_k, num_sms * 2, (a->seqlen_k + 255) / 256, 8);
      pl->mixed_form = 0;
      if (a->d == 128 && k.group <= 16) {           // grid form and count together, as for a kvcache decode (decode_choose)
        const DecodeChoice c = decode_choose((int64_t)a->b * a->h_k, a->seqlen_k, num_sms, true);
        ms = std::min(c.splits, 8); pl->mixed_form = c.form;
      }
#ifndef FA_NO_STREAMK
      if (a->b <= 512 && (int64_t)a->b * a->h_k < 4 * (int64_t)num_sms * 2 && streamk_wanted((int64_t)a->b * a->h_k, num_sms, true)) {
        // few units: the balanced-range form of the decode kernel (see the kvcache case above); one pool of piece slots
        // per launch of the half (16-row form, 32-row form)
        pl->sk_slots = num_sms * 2 + a->b * a->h_k;
        const size_t prow = (size_t)2 * pl->sk_slots * 32;
        pl->mixed_splits = 2;
        pl->mixed_lse_floats = (prow + 3) / 4 * 4;
        pl->mixed_part_off = ((pl->ws_bytes ? pl->ws_bytes : (size_t)FA_B200_WS_HEAD_BYTES) + 15) / 16 * 16;
        pl->ws_bytes = pl->mixed_part_off + (pl->mixed_lse_floats + prow * a->d) * sizeof(float);
      } else
#endif
      if (ms > 1) {
        const int pos = std::max(1, pl->decode_rows / k.group);                 // query positions a decode-half sequence can have
        const size_t prow = (size_t)ms * a->b * a->h * pos;
        pl->mixed_splits = ms;
        pl->mixed_lse_floats = (prow + 3) / 4 * 4;
        pl->mixed_part_off = ((pl->ws_bytes ? pl->ws_bytes : (size_t)FA_B200_WS_HEAD_BYTES) + 15) / 16 * 16;
        pl->ws_bytes = pl->mixed_part_off + (pl->mixed_lse_floats + prow * a->d) * sizeof(float);
      }
    }
  }
  return FA_OK;
}

static int run_plan_impl(FaFwdParams* a, int op, void* stream_);
// num_splits is an in/out field (in: the caller's request, out: the count used).  A call that FAILS must leave the
// request as it was: a caller that grows its workspace after FA_ERR_WORKSPACE and calls again with the same block would
// otherwise turn the library's own choice of the first attempt into an explicit request of the second.
static int run_plan(FaFwdParams* a, int op, void* stream_) {
  const int requested = a ? a->num_splits : 0;
  const int rc = run_plan_impl(a, op, stream_);
  if (rc != FA_OK && a) a->num_splits = requested;
  return rc;
}
static int run_plan_impl(FaFwdParams* a, int op, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  g_err[0] = 0;
  Plan pl;
  int rc = make_plan(a, op, &pl);
  if (rc != FA_OK) return rc;
  a->kernel_used = pl.kernel; a->num_splits = pl.mixed_splits > 1 ? pl.mixed_splits : pl.num_splits; a->launches = 0;
  KParams& k = pl.k;
  if (a->seqlen_q == 0 || a->b == 0) return FA_OK;
  if (pl.num_splits > 1) {
    if (a->workspace == nullptr || a->workspace_bytes < pl.ws_bytes)
      return fail(FA_ERR_WORKSPACE, "split-KV workspace too small: need %zu bytes, got %zu", pl.ws_bytes, a->workspace_bytes);
    if ((reinterpret_cast<uintptr_t>(a->workspace) & 15) != 0) return fail(FA_ERR_INVALID, "workspace must be 16-byte aligned");
    k.lse_accum = reinterpret_cast<float*>(reinterpret_cast<char*>(a->workspace) + FA_B200_WS_HEAD_BYTES);
    k.o_accum = k.lse_accum + pl.lse_floats;
    if (pl.fused_tail && !a->fused_exchange) {
      k.split_counters = pl.counters_in_head ? reinterpret_cast<int*>(a->workspace) : reinterpret_cast<int*>(k.o_accum + pl.o_floats);
      if (!(a->workspace_clean && pl.counters_in_head)) {
        cudaError_t em = cudaMemsetAsync(k.split_counters, 0, (size_t)a->b * a->h_k * sizeof(int), stream);   // a memset node, not a kernel
        if (em != cudaSuccess) return fail(FA_ERR_CUDA, "CUDA error clearing the split arrival counters: %s", cudaGetErrorString(em));
      }
    }
  }
  if (const FaFusedExchange* fx = a->fused_exchange) {
    for (int r = 0; r < fx->world; r++) { k.xchg_recv[r] = reinterpret_cast<float*>(fx->recv[r]); k.xchg_flags[r] = fx->flags[r]; }
    k.xchg_state = fx->state; k.xchg_rank = fx->rank; k.xchg_world = fx->world;
    k.xchg_max_units = fx->max_units; k.xchg_unit_elems = fx->unit_elems;
    k.split_counters = fx->split_counters;                 // zeroed once by the owner, reset by the kernel
  }
  if (k.partial && pl.num_splits == 1) {   // the single "split" IS the result: (b,h,sq,d) fp32 + lse
    k.o_accum = reinterpret_cast<float*>(a->o);
    k.lse_accum = a->softmax_lse;
  }
  cudaError_t e = cudaSuccess;
  switch (pl.kernel) {
    case FA_KERNEL_DECODE: e = launch_decode(k, pl.is_bf16, stream); break;
    case FA_KERNEL_TILE: e = launch_tile(k, pl.is_bf16, stream); break;
    case FA_KERNEL_TC5_PREFILL: {
      const bool have_list = tc5_workspace_bytes(k, a->total_q, pl.num_sms) != 0;
      if (pl.ws_bytes) {
        if (a->workspace == nullptr || a->workspace_bytes < pl.ws_bytes)
          return fail(FA_ERR_WORKSPACE, "work-list workspace too small: need %zu bytes, got %zu", pl.ws_bytes, a->workspace_bytes);
        if ((reinterpret_cast<uintptr_t>(a->workspace) & 15) != 0) return fail(FA_ERR_INVALID, "workspace must be 16-byte aligned");
      }
      if (have_list) {
        k.work_header = reinterpret_cast<int*>(reinterpret_cast<char*>(a->workspace) + FA_B200_WS_HEAD_BYTES);
        k.work_items = reinterpret_cast<int2*>(reinterpret_cast<char*>(a->workspace) + FA_B200_WS_HEAD_BYTES + 16);
      }
      a->launches = have_list ? 1 : 0;
      // ---- decode half of a mixed batch: forked onto a side stream, ahead of the prefill launches.  A batch
      // without short sequences (a plain prefill) then pays nothing on its own stream for the decode launches that
      // find no work, and in a mixed batch the tails of the two halves overlap instead of running back to back.
      ForkSet* fs = nullptr;
      bool forked2 = false;
      if (pl.decode_rows > 0) {
        // The decode kernel is instantiated for 16 and for 32 packed rows (the 32-row form holds twice the Q / O
        // fragments: fewer CTAs per SM and twice the MMA work per key).  Packed row counts are multiples of the
        // group size, so the sequences of a mixed batch are served by the 16-row form wherever they fit — single-token
        // decodes of any model with group <= 16 — and a second, 32-row launch is made only if some seqlen_q can give
        // 17..32 packed rows at all.
        const int cap = pl.decode_rows, g = k.group;
        const int lo_rows = std::min(cap, 16);
        bool second = false;
        for (int sq = 1; sq <= a->seqlen_q && sq * g <= cap; sq++) second |= sq * g > lo_rows;
        fs = get_fork_set(stream);
        cudaStream_t ds = stream;
        if (fs && cudaEventRecord(fs->fork, stream) == cudaSuccess && cudaStreamWaitEvent(fs->side, fs->fork, 0) == cudaSuccess) ds = fs->side;
        else fs = nullptr;                              // no side stream: same launches, in order, on the caller's stream
        KParams kd = k;
        if (pl.mixed_splits > 1) {
          // split decode half: partials behind the work list, arrival counters in the workspace head, merged in-kernel
          kd.num_splits = pl.mixed_splits;
          kd.part_seq_stride = std::max(1, cap / g);
          kd.lse_accum = reinterpret_cast<float*>(reinterpret_cast<char*>(a->workspace) + pl.mixed_part_off);
          kd.o_accum = kd.lse_accum + pl.mixed_lse_floats;
          kd.split_counters = reinterpret_cast<int*>(a->workspace);
          if (pl.sk_slots > 0) { kd.sk_grid = pl.num_sms * 2; kd.sk_min_per_cta = 8; kd.sk_rows = 32; }
          if (!a->workspace_clean && cudaMemsetAsync(kd.split_counters, 0, (size_t)a->b * a->h_k * sizeof(int), ds) != cudaSuccess)
            return fail(FA_ERR_CUDA, "mixed batch: clearing the split arrival counters failed");
        }
        if (g <= lo_rows) {
          kd.min_rows = 0; kd.max_rows = lo_rows;
          kd.wide_cta = (pl.mixed_form == 1 && pl.sk_slots == 0) ? 2 : 0;
          e = launch_decode(kd, pl.is_bf16, ds);
          kd.wide_cta = 0;
          a->launches += 1;
        }
        if (e == cudaSuccess && second) {
          // the 32-row form usually finds no sequence at all (it exists for 17..32 packed rows: a few-token chunk of a
          // GQA model): on a side stream of its own, so that its launch and its empty CTAs run beside the 16-row
          // launch instead of behind it
          cudaStream_t ds2 = ds;
          if (fs && g <= lo_rows && cudaStreamWaitEvent(fs->side2, fs->fork, 0) == cudaSuccess) { ds2 = fs->side2; forked2 = true; }
          kd.min_rows = lo_rows + 1; kd.max_rows = cap;
          if (pl.sk_slots > 0) {            // its own pool of piece slots
            kd.lse_accum += (size_t)pl.sk_slots * 32;
            kd.o_accum += (size_t)pl.sk_slots * 32 * a->d;
          }
          e = launch_decode(kd, pl.is_bf16, ds2);
          a->launches += 1;
          if (forked2 && cudaEventRecord(fs->join2, fs->side2) != cudaSuccess) return fail(FA_ERR_CUDA, "mixed batch: event record failed");
        }
        if (fs && cudaEventRecord(fs->join, fs->side) != cudaSuccess) return fail(FA_ERR_CUDA, "mixed batch: event record failed");

        k.min_rows = cap + 1;
      }
      if (e == cudaSuccess) e = launch_prefill_tc5(k, pl.is_bf16, a->total_q, a->total_k, pl.num_sms, stream);
      if (fs && cudaStreamWaitEvent(stream, fs->join, 0) != cudaSuccess) return fail(FA_ERR_CUDA, "mixed batch: join failed");
      if (forked2 && cudaStreamWaitEvent(stream, fs->join2, 0) != cudaSuccess) return fail(FA_ERR_CUDA, "mixed batch: join failed");
      break;
    }
    default: return fail(FA_ERR_INVALID, "unknown kernel id %d", pl.kernel);
  }
  if (e != cudaSuccess) return fail(FA_ERR_CUDA, "CUDA error in attention kernel launch: %s", cudaGetErrorString(e));
  a->launches += 1;
  if (pl.num_splits > 1 && !pl.fused_tail) {
    CombineParams c;
    c.o_part = k.o_accum; c.lse_part = k.lse_accum; c.o = a->o; c.lse = a->softmax_lse;
    c.n_splits = pl.num_splits; c.b = a->b; c.h = a->h; c.sq = a->seqlen_q; c.d = a->d;
    c.o_batch_stride = a->o_batch_stride; c.o_row_stride = a->o_row_stride; c.o_head_stride = a->o_head_stride;
    c.cu_seqlens_q = a->cu_seqlens_q; c.part_dtype = 0; c.out_dtype = k.partial ? 0 : (pl.is_bf16 ? 1 : 2);
    c.part_split_stride = (int64_t)a->b * a->h * a->seqlen_q * a->d;
    c.lse_split_stride = (int64_t)a->b * a->h * a->seqlen_q;
    c.empty_neg_inf = k.partial;
    e = launch_combine(c, stream);
    if (e != cudaSuccess) return fail(FA_ERR_CUDA, "CUDA error in combine kernel launch: %s", cudaGetErrorString(e));
    a->launches = 2;
  }
  return FA_OK;
}

extern "C" size_t fa_b200_workspace_bytes(const FaFwdParams* p, int op) {
  Plan pl;
  if (op < 0 || op > 2 || make_plan(p, op, &pl) != FA_OK) return (size_t)-1;
  return pl.ws_bytes ? pl.ws_bytes + 16 : 0;
}

// mha_fwd (flash_api.cpp:240-420): dense (b,s,h,d) tensors.
extern "C" int fa_b200_fwd(FaFwdParams* p, void* stream) {
  if (p && (p->block_table || p->cache_batch_idx || p->cu_seqlens_k || p->seqused_k))
    return fail(FA_ERR_INVALID, "fwd takes dense tensors only (no block_table / cache_batch_idx / seqlens)");
  return run_plan(p, OP_FWD, stream);
}

// mha_varlen_fwd (flash_api.cpp:422-669): packed (total,h,d) q with cu_seqlens; k/v packed or paged.
extern "C" int fa_b200_varlen_fwd(FaFwdParams* p, void* stream) {
  if (p && p->cache_batch_idx) return fail(FA_ERR_INVALID, "varlen_fwd does not take cache_batch_idx");
  return run_plan(p, OP_VARLEN, stream);
}

// mha_fwd_kvcache (flash_api.cpp:671-937): q (b,sq,h,d) against a (paged) cache with per-sequence
// lengths in cu_seqlens_k (not cumulative, :859-867).
extern "C" int fa_b200_fwd_kvcache(FaFwdParams* p, void* stream) {
  return run_plan(p, OP_KVCACHE, stream);
}

extern "C" int fa_b200_combine(const void* o_part, const float* lse_part, void* o, float* lse, int32_t n,
                               int64_t rows, int32_t d, int64_t part_stride, int64_t lse_stride,
                               int32_t part_dtype, int32_t out_dtype, void* stream) {
  g_err[0] = 0;
  if (!o_part || !lse_part || !o) return fail(FA_ERR_INVALID, "combine: NULL pointer");
  if (n < 1 || n > 128) return fail(FA_ERR_INVALID, "combine: n must be in [1,128]");
  if (d % 4 != 0 || d <= 0) return fail(FA_ERR_INVALID, "combine: d must be a positive multiple of 4");
  if (rows > 0x7fffffff || rows < 0) return fail(FA_ERR_INVALID, "combine: bad row count");
  if (part_stride < rows * d || lse_stride < rows) return fail(FA_ERR_INVALID, "combine: partial strides smaller than one partial");
  const int esz = part_dtype == 0 ? 4 : 2;
  if ((reinterpret_cast<uintptr_t>(o_part) % (4 * esz)) != 0 || (part_stride * esz) % (4 * esz) != 0)
    return fail(FA_ERR_INVALID, "combine: partials must be aligned to 4 elements");
  CombineParams c;
  c.o_part = o_part; c.lse_part = lse_part; c.o = o; c.lse = lse;
  c.n_splits = n; c.b = 1; c.h = 1; c.sq = (int)rows; c.d = d;
  c.o_batch_stride = 0; c.o_row_stride = d; c.o_head_stride = 0; c.cu_seqlens_q = nullptr;
  c.part_dtype = part_dtype; c.out_dtype = out_dtype;
  c.part_split_stride = part_stride; c.lse_split_stride = lse_stride; c.empty_neg_inf = 0;
  cudaError_t e = launch_combine(c, (cudaStream_t)stream);
  if (e != cudaSuccess) return fail(FA_ERR_CUDA, "CUDA error in combine kernel launch: %s", cudaGetErrorString(e));
  return FA_OK;
}

// Append (and rotate) the new K/V rows of a fwd_kvcache call — the first half of the reference's
// fused append path (flash_fwd_kernel.h:672-824).  Takes the same parameter block as
// fa_b200_fwd_kvcache with knew/vnew (and optionally rotary_cos/sin) set; if rotary is given,
// `workspace` must hold b*seqlen_q*h*d elements and receives the rotated q, which the caller then
// passes as q to fa_b200_fwd_kvcache together with seqlen_knew.
extern "C" int fa_b200_append_kv(FaFwdParams* a, void* stream) {
  g_err[0] = 0;
  FA_CHECK(a && a->knew && a->vnew, "append_kv: knew and vnew must be given");
  FA_CHECK(a->cu_seqlens_k != nullptr, "seqlens_k must be passed if key/value are passed");
  FA_CHECK(a->seqlen_knew > 0, "append_kv: seqlen_knew must be positive");
  FA_CHECK(a->d % 8 == 0, "append_kv: head dimension must be a multiple of 8");
  const bool rotary = a->rotary_cos != nullptr;
  if (rotary) {
    FA_CHECK(a->rotary_sin != nullptr, "If rotary cos is provided, rotary sin must also be provided");
    FA_CHECK(a->rotary_dim <= a->d, "rotary_dim must be <= headdim");
    FA_CHECK(a->rotary_dim % 16 == 0, "Only rotary dimensions divisible by 16 are currently supported");
    FA_CHECK(a->workspace && a->workspace_bytes >= (size_t)a->b * a->seqlen_q * a->h * a->d * 2,
             "append_kv: workspace for the rotated q is too small");
  }
  if (a->block_table) FA_CHECK(a->page_block_size > 0 && a->page_block_size % 16 == 0, "Paged KV cache block size must be divisible by 16");
  // same mask normalisation as the attention call decides whether q rows rotate by position
  int wl = a->window_left, wr = a->window_right;
  bool causal = a->is_causal != 0;
  if (wl >= a->seqlen_k) wl = -1;
  if (wr >= a->seqlen_k) wr = -1;
  if (a->seqlen_q == 1 && !a->alibi_slopes) causal = false;
  if (causal) wr = 0;
  AppendParams p;
  memset(&p, 0, sizeof p);
  p.knew = a->knew; p.vnew = a->vnew; p.kcache = const_cast<void*>(a->k); p.vcache = const_cast<void*>(a->v);
  p.q = a->q; p.q_rot = rotary ? a->workspace : nullptr;
  p.knew_batch_stride = a->knew_batch_stride; p.knew_row_stride = a->knew_row_stride; p.knew_head_stride = a->knew_head_stride;
  p.vnew_batch_stride = a->vnew_batch_stride; p.vnew_row_stride = a->vnew_row_stride; p.vnew_head_stride = a->vnew_head_stride;
  p.k_batch_stride = a->k_batch_stride; p.k_row_stride = a->k_row_stride; p.k_head_stride = a->k_head_stride;
  p.v_batch_stride = a->v_batch_stride; p.v_row_stride = a->v_row_stride; p.v_head_stride = a->v_head_stride;
  p.q_batch_stride = a->q_batch_stride; p.q_row_stride = a->q_row_stride; p.q_head_stride = a->q_head_stride;
  p.b = a->b; p.h = a->h; p.h_k = a->h_k; p.d = a->d; p.seqlen_q = a->seqlen_q; p.seqlen_k = a->seqlen_k;
  p.seqlen_knew = a->seqlen_knew;
  p.cache_seqlens = a->cu_seqlens_k; p.cache_batch_idx = a->cache_batch_idx;
  p.block_table = a->block_table; p.block_table_batch_stride = a->block_table_batch_stride;
  p.page_block_size = a->page_block_size; p.max_blocks_per_seq = a->max_blocks_per_seq;
  p.rotary_cos = a->rotary_cos; p.rotary_sin = a->rotary_sin;
  p.rotary_dim = rotary ? a->rotary_dim : 0; p.seqlen_ro = a->seqlen_ro; p.is_rotary_interleaved = a->is_rotary_interleaved;
  p.rot_by_qpos = (wr >= 0 || wl >= 0) ? 1 : 0;   // causal or local after normalisation
  cudaError_t e = launch_append(p, a->is_bf16 != 0, (cudaStream_t)stream);
  if (e != cudaSuccess) return fail(FA_ERR_CUDA, "CUDA error in append kernel launch: %s", cudaGetErrorString(e));
  return FA_OK;
}

// Cache writer for the caller's side of the path (vLLM: reshape_and_cache_flash): see fa_append.cu.
extern "C" int fa_b200_reshape_and_cache(const void* key, const void* value, void* kcache, void* vcache, const int64_t* slot_mapping,
                                         int num_tokens, int h_k, int d, int page_block_size, int64_t num_blocks,
                                         int64_t key_token_stride, int64_t key_head_stride, int64_t value_token_stride, int64_t value_head_stride,
                                         int64_t k_batch_stride, int64_t k_row_stride, int64_t k_head_stride,
                                         int64_t v_batch_stride, int64_t v_row_stride, int64_t v_head_stride, void* stream) {
  g_err[0] = 0;
  FA_CHECK(num_tokens >= 0 && h_k > 0 && d > 0 && d % 8 == 0, "reshape_and_cache: bad shape (head dimension must be a multiple of 8)");
  if (num_tokens == 0) return FA_OK;
  FA_CHECK(key && value && kcache && vcache && slot_mapping, "reshape_and_cache: NULL pointer");
  FA_CHECK(page_block_size > 0 && num_blocks > 0, "reshape_and_cache: bad page_block_size / num_blocks");
  FA_CHECK((reinterpret_cast<uintptr_t>(key) | reinterpret_cast<uintptr_t>(value) | reinterpret_cast<uintptr_t>(kcache) | reinterpret_cast<uintptr_t>(vcache)) % 16 == 0 &&
           (key_token_stride | key_head_stride | value_token_stride | value_head_stride | k_batch_stride | k_row_stride | k_head_stride |
            v_batch_stride | v_row_stride | v_head_stride) % 8 == 0, "reshape_and_cache: tensors and strides must be 16-byte aligned");
  CacheWriteParams a;
  a.key = key; a.value = value; a.kcache = kcache; a.vcache = vcache; a.slot_mapping = slot_mapping;
  a.key_token_stride = key_token_stride; a.key_head_stride = key_head_stride; a.value_token_stride = value_token_stride; a.value_head_stride = value_head_stride;
  a.k_batch_stride = k_batch_stride; a.k_row_stride = k_row_stride; a.k_head_stride = k_head_stride;
  a.v_batch_stride = v_batch_stride; a.v_row_stride = v_row_stride; a.v_head_stride = v_head_stride;
  a.num_tokens = num_tokens; a.h_k = h_k; a.d = d; a.page = page_block_size;
  cudaError_t e = launch_reshape_and_cache(a, (cudaStream_t)stream);
  if (e != cudaSuccess) return fail(FA_ERR_CUDA, "CUDA error in cache-write kernel launch: %s", cudaGetErrorString(e));
  return FA_OK;
}

// Will fa_b200_fwd_kvcache append p->knew / p->vnew (and apply rotary) inside its own kernel?  1 = yes (pass them to
// fwd_kvcache directly: one launch), 0 = no (run fa_b200_append_kv first), negative = invalid arguments.
extern "C" int fa_b200_fused_append_supported(const FaFwdParams* a) {
  g_err[0] = 0;
  FA_CHECK(a != nullptr, "params is NULL");
  FaFwdParams t = *a;
  t.knew = t.vnew = t.rotary_cos = t.rotary_sin = nullptr;       // plan the attention call alone
  Plan pl;
  const int rc = make_plan(&t, OP_KVCACHE, &pl);
  if (rc != FA_OK) return rc;
  return pl.kernel == FA_KERNEL_DECODE && !pl.k.is_local && a->kv_shard_len == 0 ? 1 : 0;
}

// Sequence-sharded decode: push + wait + LSE-combine across the GPUs of one box (csrc/fa_exchange.cu).
extern "C" int fa_b200_exchange_combine(const FaExchangeParams* a, void* stream) {
  g_err[0] = 0;
  FA_CHECK(a != nullptr, "params is NULL");
  FA_CHECK(a->world >= 1 && a->world <= FA_MAX_PEERS && a->rank >= 0 && a->rank < a->world, "bad rank/world");
  FA_CHECK(a->b > 0 && a->h > 0 && a->seqlen_q > 0 && a->d > 0 && a->d % 4 == 0, "bad shape");
  FA_CHECK(a->record_elems % 4 == 0 && a->record_elems >= (int64_t)a->b * a->h * a->seqlen_q * (a->d + 1), "record_elems too small or not a multiple of 4");
  FA_CHECK(a->local_record && a->state && a->o, "NULL pointer");
  FA_CHECK((reinterpret_cast<uintptr_t>(a->local_record) & 15) == 0, "local_record must be 16-byte aligned");
  FA_CHECK(a->out_dtype >= 0 && a->out_dtype <= 2, "bad out_dtype");
  for (int r = 0; r < a->world; r++)
    FA_CHECK(a->recv[r] && a->flags[r] && (reinterpret_cast<uintptr_t>(a->recv[r]) & 15) == 0, "recv/flags pointer missing or misaligned");
  FA_CHECK(a->o_row_stride % 4 == 0 && a->o_head_stride % 4 == 0 && a->o_batch_stride % 4 == 0, "output strides must be multiples of 4 elements");
  cudaError_t e = launch_exchange_combine(*a, (cudaStream_t)stream);
  if (e != cudaSuccess) return fail(FA_ERR_CUDA, "CUDA error in exchange kernel launch: %s", cudaGetErrorString(e));
  return FA_OK;
}
extern "C" size_t fa_b200_exchange_params_size(void) { return sizeof(FaExchangeParams); }
extern "C" size_t fa_b200_fused_exchange_size(void) { return sizeof(FaFusedExchange); }

// ---------------------------------------------------------------------------------------------
// fa_b200_fwd_kvcache_static_scratch — fwd_kvcache for callers that cannot allocate (CUDA-graph
// capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Same call as fa_b200_fwd_kvcache, but the split-KV
// partials (lse_accum | o_accum) that fa_decode_kernel / fa_tile_kernel write and fa_combine_kernel
// merges live in the file-scope device array below instead of a caller-allocated workspace.
//   * Size: FA_B200_STATIC_SCRATCH_BYTES (4 MiB) — every GPU that loads the library holds it until exit.
//     A call whose plan needs more is rejected (FA_ERR_WORKSPACE) before anything is launched.
//   * The array's NAME is only a host-side stand-in in host code: its device address is fetched with
//     cudaGetSymbolAddress once per device (that device current) and that pointer goes to the kernels.
//   * Every call shares the array and finds the previous call's contents in it.  What has to start at
//     zero are the arrival counters of the in-kernel split merge (workspace head): the array is zero when
//     the module loads and the kernel's last CTA per unit puts its counter back to zero.  The partials
//     need no clearing: every CTA of the split kernel writes its (split, row) slice, empty splits included
//     (O = 0, lse = -inf), before the unit's last CTA reads them, and the prefill work
//     list (when a many-row call lands on the tcgen05 kernel) is rebuilt by tc5_schedule_kernel, which
//     also resets its draw counter.  Calls that OVERLAP on the GPU (different streams, or graphs
//     replayed concurrently) race on the array: serialise them, or use fa_b200_fwd_kvcache with
//     separate workspaces.
//   * Calls no cudaMalloc / cudaFree / allocating library routine.
static __device__ __align__(256) unsigned char g_static_scratch[FA_B200_STATIC_SCRATCH_BYTES];

static int static_scratch_ptr(void** out) {
  static void* ptrs[64] = {nullptr};           // per device; a racing first call stores the same value twice
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail(FA_ERR_CUDA, "static scratch: cudaGetDevice failed: %s", cudaGetErrorString(e));
  if (dev < 0 || dev >= 64) return fail(FA_ERR_UNSUPPORTED, "static scratch: device index %d out of range", dev);
  if (!ptrs[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || p == nullptr)
      return fail(FA_ERR_CUDA, "static scratch: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return FA_OK;
}

extern "C" int fa_b200_fwd_kvcache_static_scratch(FaFwdParams* a, void* stream) {
  g_err[0] = 0;
  FA_CHECK(a != nullptr, "params is NULL");
  FA_CHECK(a->workspace == nullptr && a->workspace_bytes == 0,
           "fwd_kvcache_static_scratch takes no workspace: the scratch is the library's static device array");
  Plan pl;
  int rc = make_plan(a, OP_KVCACHE, &pl);
  if (rc != FA_OK) return rc;
  const size_t need = pl.ws_bytes ? pl.ws_bytes + 16 : 0;     // + alignment slack, as fa_b200_workspace_bytes reports
  if (need > (size_t)FA_B200_STATIC_SCRATCH_BYTES)
    return fail(FA_ERR_WORKSPACE, "fwd_kvcache_static_scratch: this call needs %zu bytes of scratch, the static maximum is %zu "
                "(use fa_b200_fwd_kvcache with a caller-allocated workspace, or fewer splits)",
                need, (size_t)FA_B200_STATIC_SCRATCH_BYTES);
  void* ptr = nullptr;
  if (need) {
    rc = static_scratch_ptr(&ptr);
    if (rc != FA_OK) return rc;
  }
  a->workspace = ptr;
  a->workspace_bytes = need ? (size_t)FA_B200_STATIC_SCRATCH_BYTES : 0;
  const int clean_in = a->workspace_clean;
  a->workspace_clean = 1;      // the static array is zero when the module loads and the arrival counters reset themselves
  rc = run_plan(a, OP_KVCACHE, stream);
  a->workspace = nullptr; a->workspace_bytes = 0; a->workspace_clean = clean_in;   // the caller's block stays as it was passed
  return rc;
}
