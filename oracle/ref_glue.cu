// TEST INFRASTRUCTURE — not part of the product path.
//
// Thin C-ABI launcher around the REFERENCE's own FA2 kernels, compiled for sm_100 from the
// sources where they lie under /root/reference (see oracle/Makefile; nothing is copied into
// this repo).  It exists so the parity tests can compare libfa_b200.so with the reference
// kernel on the same inputs (north_star: "max-abs error no worse than 2x the reference
// kernel's own error") and so bench.py can time "the reference kernel on B200".
//
// The reference's pybind glue (flash_api.cpp) cannot be used as is: its runtime gate rejects
// compute capability 10.x (flash_api.cpp:256-258,445-447,694-696).  This file is our own
// launcher: it fills the reference's Flash_fwd_params (src/flash.h:51-141) from our
// FaFwdParams following set_params_fprop (flash_api.cpp:21-138), the GQA decode swap
// (flash_api.cpp:298-304,500-507,749-755), set_params_splitkv (:194-219) and run_mha_fwd
// (:140-150), then calls the reference's run_mha_fwd_<T,d> / run_mha_fwd_splitkv_dispatch<T,d>
// template instantiations (src/flash_fwd_hdim*_sm80.cu, src/flash_fwd_split_hdim*_sm80.cu).
// Only d in {64,128} x {bf16,fp16} are built.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <algorithm>
#include <cuda_runtime.h>
#include <cutlass/numeric_types.h>
#include "flash.h"          // reference header, via -I/root/reference/csrc/flash_attn/src
#include "fa_b200.h"

template <typename T, int Headdim> void run_mha_fwd_(Flash_fwd_params& params, cudaStream_t stream);
template <typename T, int Headdim> void run_mha_fwd_splitkv_dispatch(Flash_fwd_params& params, cudaStream_t stream);

static thread_local char g_err[512] = "";

static int round_up(int x, int m) { return (x + m - 1) / m * m; }

// flash_api.cpp:158-192 restated (kept separate from the product's copy on purpose).
static int ref_splits_heuristic(int work, int num_sms, int n_blocks, int max_splits) {
  if (work >= 0.8f * num_sms) return 1;
  max_splits = std::min({max_splits, num_sms, n_blocks});
  auto cdiv = [](int a, int b) { return (a + b - 1) / b; };
  auto eligible = [&](int s) { return s == 1 || cdiv(n_blocks, s) != cdiv(n_blocks, s - 1); };
  float best = 0.f;
  float eff[129];
  for (int s = 1; s <= max_splits; s++) {
    eff[s] = 0.f;
    if (!eligible(s)) continue;
    float waves = float(work * s) / num_sms;
    eff[s] = waves / std::ceil(waves);
    best = std::max(best, eff[s]);
  }
  for (int s = 1; s <= max_splits; s++)
    if (eligible(s) && eff[s] >= 0.85f * best) return s;
  return 1;
}

extern "C" const char* ref_fa2_last_error(void) { return g_err; }

// Workspace needed for op (0 fwd, 1 varlen, 2 kvcache): fp32 (S,b,h,sq) lse + (S,b,h,sq,d) out.
// Always reports the 128-split worst case so the caller need not know the heuristic.
extern "C" size_t ref_fa2_workspace_bytes(const FaFwdParams* p) {
  size_t rows = (size_t)p->b * p->h * std::max(p->seqlen_q, 1);
  return 128 * rows * (size_t)(round_up(p->d, 32) + 1) * sizeof(float) + 16;
}

// op: 0 = mha_fwd, 1 = mha_varlen_fwd, 2 = mha_fwd_kvcache.
extern "C" int ref_fa2_run(FaFwdParams* a, int op, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  g_err[0] = 0;
  if (a->d != 64 && a->d != 128) { snprintf(g_err, sizeof g_err, "ref build has d in {64,128} only"); return FA_ERR_UNSUPPORTED; }
  if (a->knew || a->rotary_cos) { snprintf(g_err, sizeof g_err, "ref glue: append-KV not wired"); return FA_ERR_UNSUPPORTED; }

  int b = a->b, h = a->h, h_k = a->h_k, d = a->d;
  int seqlen_q = a->seqlen_q, seqlen_k = a->seqlen_k;
  int wl = a->window_left, wr = a->window_right;
  bool causal = a->is_causal != 0;
  const bool paged = a->block_table != nullptr;

  if (wl >= seqlen_k) wl = -1;             // flash_api.cpp:289-290
  if (wr >= seqlen_k) wr = -1;
  if (seqlen_q == 1 && !a->alibi_slopes) causal = false;   // :293,:493,:744
  if (causal) wr = 0;                       // :294

  const bool swapped = seqlen_q == 1 && h > h_k && wl < 0 && wr < 0 && d % 8 == 0 && !a->alibi_slopes;

  Flash_fwd_params P;
  memset(&P, 0, sizeof P);
  P.is_bf16 = a->is_bf16 != 0;
  P.q_ptr = const_cast<void*>(a->q); P.k_ptr = const_cast<void*>(a->k); P.v_ptr = const_cast<void*>(a->v);
  P.o_ptr = a->o;
  P.q_row_stride = a->q_row_stride; P.k_row_stride = a->k_row_stride; P.v_row_stride = a->v_row_stride;
  P.q_head_stride = a->q_head_stride; P.k_head_stride = a->k_head_stride; P.v_head_stride = a->v_head_stride;
  P.o_row_stride = a->o_row_stride; P.o_head_stride = a->o_head_stride;
  P.q_batch_stride = a->q_batch_stride; P.k_batch_stride = a->k_batch_stride; P.v_batch_stride = a->v_batch_stride;
  P.o_batch_stride = a->o_batch_stride;
  P.cu_seqlens_q = const_cast<int*>(a->cu_seqlens_q);
  P.cu_seqlens_k = const_cast<int*>(a->cu_seqlens_k);
  P.seqused_k = const_cast<int*>(a->seqused_k);
  P.is_seqlens_k_cumulative = a->is_seqlens_k_cumulative != 0;

  if (swapped) {
    // q (b,1,h_k*g,d) -> (b,g,h_k,d): heads of one group become query rows.
    const int g = h / h_k;
    P.q_row_stride = a->q_head_stride;  P.q_head_stride = a->q_head_stride * g;
    P.o_row_stride = a->o_head_stride;  P.o_head_stride = a->o_head_stride * g;
    if (op == 1) {  // varlen pure-decode: q is (b,h,d) with row stride as batch stride
      P.q_batch_stride = a->q_row_stride; P.o_batch_stride = a->o_row_stride;
      P.cu_seqlens_q = nullptr;
    }
    seqlen_q = g; h = h_k;
  }

  P.b = b; P.h = h; P.h_k = h_k; P.h_h_k_ratio = h / h_k;
  P.seqlen_q = seqlen_q; P.seqlen_k = seqlen_k;
  P.seqlen_q_rounded = round_up(seqlen_q, 128); P.seqlen_k_rounded = round_up(seqlen_k, 128);
  P.d = d; P.d_rounded = round_up(d, 32);
  P.scale_softmax = a->softmax_scale;
  P.scale_softmax_log2 = a->softmax_scale * float(M_LOG2E);
  P.p_dropout = 1.f; P.p_dropout_in_uint8_t = 255; P.rp_dropout = 1.f;
  P.scale_softmax_rp_dropout = a->softmax_scale;
  P.is_causal = wl < 0 && wr == 0;
  if (wl < 0 && wr >= 0) wl = seqlen_k;
  if (wl >= 0 && wr < 0) wr = seqlen_k;
  P.window_size_left = wl; P.window_size_right = wr;
  P.softmax_lse_ptr = a->softmax_lse;
  P.alibi_slopes_ptr = const_cast<float*>(a->alibi_slopes);
  P.alibi_slopes_batch_stride = a->alibi_slopes_batch_stride;
  P.block_table = const_cast<int*>(a->block_table);
  P.block_table_batch_stride = a->block_table_batch_stride;
  P.page_block_size = a->page_block_size;
  P.cache_batch_idx = const_cast<int*>(a->cache_batch_idx);

  int dev = 0, num_sms = a->num_sms;
  if (num_sms <= 0) { cudaGetDevice(&dev); cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev); }

  // Split choice: mha_fwd always consults the heuristic (:375-377), varlen only when
  // swapped (:618-623), kvcache with the caller's num_splits (:903-905).
  P.num_splits = 1;
  bool consult = op == 0 || (op == 1 && swapped) || op == 2;
  if (consult) {
    int req = op == 2 ? a->num_splits : 0;
    const int block_n = d <= 64 ? 256 : (d <= 128 ? 128 : 64);
    const int n_blocks = (seqlen_k + block_n - 1) / block_n;
    const int m_blocks = (seqlen_q + 63) / 64;
    P.num_splits = req >= 1 ? req : ref_splits_heuristic(b * h * m_blocks, num_sms * 2, n_blocks, 128);
    if (P.num_splits > 128) { snprintf(g_err, sizeof g_err, "num_splits > 128 not supported"); return FA_ERR_INVALID; }
    if (P.num_splits > 1) {
      size_t rows = (size_t)b * h * seqlen_q;
      size_t need = (size_t)P.num_splits * rows * (P.d_rounded + 1) * sizeof(float) + 16;
      if (!a->workspace || a->workspace_bytes < need) { snprintf(g_err, sizeof g_err, "workspace too small"); return FA_ERR_WORKSPACE; }
      P.softmax_lseaccum_ptr = a->workspace;
      P.oaccum_ptr = (float*)a->workspace + ((size_t)P.num_splits * rows + 3) / 4 * 4;   // the combine kernel reads 16-byte pieces
    }
  }
  a->num_splits = P.num_splits;
  const bool force_split = (op == 1 && paged) || (op == 2 && (paged || a->cache_batch_idx));

  if (seqlen_k > 0) {
    const bool split_path = P.num_splits > 1 || force_split;
#define REF_LAUNCH(T, D) do { if (split_path) run_mha_fwd_splitkv_dispatch<T, D>(P, stream); else run_mha_fwd_<T, D>(P, stream); } while (0)
    try {
      if (P.is_bf16) { if (d == 64) REF_LAUNCH(cutlass::bfloat16_t, 64); else REF_LAUNCH(cutlass::bfloat16_t, 128); }
      else           { if (d == 64) REF_LAUNCH(cutlass::half_t, 64);     else REF_LAUNCH(cutlass::half_t, 128); }
    } catch (const std::exception& e) {
      snprintf(g_err, sizeof g_err, "%s", e.what()); return FA_ERR_CUDA;
    }
#undef REF_LAUNCH
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { snprintf(g_err, sizeof g_err, "CUDA: %s", cudaGetErrorString(e)); return FA_ERR_CUDA; }
  return FA_OK;
}
