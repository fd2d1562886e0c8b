// pybind11 / ATen host binding over the C ABI of libfa_b200.so — the compiled form of the reference's extension
// module (PYBIND11_MODULE in csrc/flash_attn/flash_api.cpp:939-946): fwd / varlen_fwd / fwd_kvcache with the
// positional argument lists and return lists of mha_fwd (:240-252), mha_varlen_fwd (:422-441) and
// mha_fwd_kvcache (:671-690).  INTEGRATION.md "Option B2" made real: what a maintainer who keeps the pybind11
// boundary links against -lfa_b200.  Only tensor plumbing lives here (checks, out / lse / workspace allocation,
// strides, current stream); masks, kernel choice, splits and launches are in the library (fa_api.cu).
// The same logic exists in Python over ctypes (vllm_flash_attn/vllm_flash_attn_2_cuda.py), which also carries the
// sharding extensions; this file is the low-overhead path for the three plain ops.
#include <torch/extension.h>
#include <ATen/cuda/CUDAContext.h>
#include <cuda_runtime_api.h>
#include <c10/cuda/CUDAGuard.h>

#include <thread>
#include <unordered_map>
#include <vector>

#include "fa_b200.h"

namespace {

struct CallInfo { int kernel_used = 0, num_splits = 0, launches = 0; };
thread_local CallInfo g_info;          // diagnostics of the last call on this host thread

inline void ok(int rc) { TORCH_CHECK(rc == FA_OK, fa_b200_last_error()); }

void check_qkv(const at::Tensor& q, const at::Tensor& k, const at::Tensor& v) {
  TORCH_CHECK(q.dtype() == torch::kFloat16 || q.dtype() == torch::kBFloat16, "FlashAttention only support fp16 and bf16 data type");
  TORCH_CHECK(k.dtype() == q.dtype(), "query and key must have the same dtype");
  TORCH_CHECK(v.dtype() == q.dtype(), "query and value must have the same dtype");
  TORCH_CHECK(q.is_cuda(), "q must be on CUDA"); TORCH_CHECK(k.is_cuda(), "k must be on CUDA"); TORCH_CHECK(v.is_cuda(), "v must be on CUDA");
  TORCH_CHECK(q.stride(-1) == 1 && k.stride(-1) == 1 && v.stride(-1) == 1, "Input tensor must have contiguous last dimension");
}

void check_i32(const at::Tensor& t, const char* name, const at::Device& dev) {
  TORCH_CHECK(t.dtype() == torch::kInt32, name, " must have dtype int32");
  TORCH_CHECK(t.is_cuda() && t.device() == dev, name, " must be on the same CUDA device as q");
}

void set_alibi(FaFwdParams& P, const c10::optional<at::Tensor>& a_, int64_t b, int64_t h, const at::Device& dev) {
  if (!a_.has_value()) return;
  const at::Tensor& a = *a_;
  TORCH_CHECK(a.dtype() == torch::kFloat32, "ALiBi slopes must have dtype fp32");
  TORCH_CHECK(a.is_cuda() && a.device() == dev, "ALiBi slopes must be on the same device");
  TORCH_CHECK(a.stride(-1) == 1, "ALiBi slopes tensor must have contiguous last dimension");
  TORCH_CHECK((a.dim() == 1 && a.size(0) == h) || (a.dim() == 2 && a.size(0) == b && a.size(1) == h),
              "ALiBi slopes must have shape (nheads,) or (batch, nheads)");
  P.alibi_slopes = a.data_ptr<float>();
  P.alibi_slopes_batch_stride = a.dim() == 2 ? a.stride(0) : 0;
}

at::Tensor pad_head(const at::Tensor& x, int64_t d) {
  return x.size(-1) == d ? x : torch::nn::functional::pad(x, torch::nn::functional::PadFuncOptions({0, d - x.size(-1)}));
}

// Persistent zero-initialised workspace per (device, stream) of THIS host thread — the library keeps the arrival counters
// at its head at zero between calls (workspace_clean), so a split decode is one launch.  Under stream capture a private
// zeroed buffer is used instead so that the capture owns it.  The map is leaked on purpose (no CUDA frees at exit).
struct WsKey { int dev; cudaStream_t s; bool operator==(const WsKey& o) const { return dev == o.dev && s == o.s; } };
struct WsHash { size_t operator()(const WsKey& k) const { return std::hash<void*>()((void*)k.s) ^ (size_t)k.dev * 0x9e3779b97f4a7c15ull; } };
std::unordered_map<WsKey, at::Tensor, WsHash>& ws_map() {
  thread_local auto* m = new std::unordered_map<WsKey, at::Tensor, WsHash>();
  return *m;
}

using Fn = int (*)(FaFwdParams*, void*);
void run(FaFwdParams& P, int op, Fn fn, const at::Tensor& like) {
  auto stream = at::cuda::getCurrentCUDAStream(like.get_device());
  auto opts = like.options().dtype(torch::kUInt8);
  P.workspace_clean = 1;
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(stream.stream(), &cap) != cudaSuccess) { (void)cudaGetLastError(); cap = cudaStreamCaptureStatusNone; }
  if (cap != cudaStreamCaptureStatusNone) {
    const size_t need = fa_b200_workspace_bytes(&P, op);
    TORCH_CHECK(need != (size_t)-1, fa_b200_last_error());
    at::Tensor ws = torch::zeros({(int64_t)std::max<size_t>(need, 16)}, opts);
    P.workspace = ws.data_ptr(); P.workspace_bytes = (size_t)ws.numel();
    ok(fn(&P, stream.stream()));
  } else {
    auto& m = ws_map();
    const WsKey key{like.get_device(), stream.stream()};
    auto it = m.find(key);
    if (it != m.end()) { P.workspace = it->second.data_ptr(); P.workspace_bytes = (size_t)it->second.numel(); }
    int rc = fn(&P, stream.stream());
    if (rc == FA_ERR_WORKSPACE) {
      const size_t need = fa_b200_workspace_bytes(&P, op);
      TORCH_CHECK(need != (size_t)-1, fa_b200_last_error());
      at::Tensor ws = torch::zeros({(int64_t)std::max<size_t>(2 * need, (size_t)1 << 20)}, opts);
      m[key] = ws;
      P.workspace = ws.data_ptr(); P.workspace_bytes = (size_t)ws.numel();
      rc = fn(&P, stream.stream());
    }
    ok(rc);
  }
  g_info.kernel_used = P.kernel_used; g_info.num_splits = P.num_splits; g_info.launches = P.launches;
}

// `p` (empty: no dropout) and `rng_state` outputs of fwd / varlen_fwd (flash_api.cpp:384-392,409): never written
// (one pair per (dtype, device) and host thread instead of two allocator round trips per call; leaked like the workspaces)
std::pair<at::Tensor, at::Tensor> placeholders(const at::Tensor& q) {
  thread_local auto* cache = new std::unordered_map<int, std::pair<at::Tensor, at::Tensor>>();
  const int key = q.get_device() * 64 + (int)q.scalar_type();
  auto it = cache->find(key);
  if (it == cache->end())
    it = cache->emplace(key, std::make_pair(torch::empty({0}, q.options()), torch::zeros({2}, q.options().dtype(torch::kInt64)))).first;
  return it->second;
}

at::Tensor make_out(const c10::optional<at::Tensor>& out_, const at::Tensor& q, const at::Tensor& q_p, int64_t d, int64_t d_og) {
  if (out_.has_value()) {
    const at::Tensor& o = *out_;
    TORCH_CHECK(o.dtype() == q.dtype(), "Output must have the same dtype as inputs");
    TORCH_CHECK(o.is_cuda() && o.device() == q.device(), "Output must be on the same device");
    TORCH_CHECK(o.stride(-1) == 1, "Output tensor must have contiguous last dimension");
    TORCH_CHECK(o.sizes() == q.sizes(), "out has the wrong shape");
    return d == d_og ? o : torch::empty_like(q_p);
  }
  return torch::empty_like(q_p);
}

std::vector<at::Tensor> finish(at::Tensor out, const c10::optional<at::Tensor>& out_, const at::Tensor& q_p, const at::Tensor& k_p,
                               const at::Tensor& v_p, const at::Tensor& lse, const at::Tensor& q, int64_t d, int64_t d_og) {
  at::Tensor out_padded = out;
  if (d != d_og) {
    out = out.slice(-1, 0, d_og);
    if (out_.has_value()) { out_->copy_(out); out = *out_; }
  }
  auto ph = placeholders(q);
  return {out, q_p, k_p, v_p, out_padded, lse, ph.first, ph.second};
}

// mha_fwd (flash_api.cpp:240-420).  q (b,sq,h,d); k,v (b,sk,hk,d).
// Returns {out, q_padded, k_padded, v_padded, out_padded, softmax_lse, p, rng_state}.
std::vector<at::Tensor> fwd(at::Tensor& q, const at::Tensor& k, const at::Tensor& v, c10::optional<at::Tensor>& out_,
                            c10::optional<at::Tensor>& alibi_slopes_, const double p_dropout, const double softmax_scale,
                            bool is_causal, int64_t window_size_left, int64_t window_size_right, const bool return_softmax,
                            c10::optional<at::Generator> gen_) {
  check_qkv(q, k, v);
  TORCH_CHECK(p_dropout == 0.0, "This flash attention build does not support dropout.");
  TORCH_CHECK(q.dim() == 4 && k.dim() == 4 && v.dim() == 4, "q, k, v must be (batch, seqlen, nheads, headdim)");
  const int64_t b = q.size(0), sq = q.size(1), h = q.size(2), d_og = q.size(3), sk = k.size(1), hk = k.size(2);
  TORCH_CHECK(b > 0, "batch size must be postive");
  TORCH_CHECK(d_og <= 256, "FlashAttention forward only supports head dimension at most 256");
  TORCH_CHECK(h % hk == 0, "Number of heads in key/value must divide number of heads in query");
  TORCH_CHECK(k.size(0) == b && k.size(3) == d_og && v.sizes() == k.sizes(), "k/v shape mismatch");
  const int64_t d = (d_og + 7) / 8 * 8;
  at::Tensor q_p = pad_head(q, d), k_p = pad_head(k, d), v_p = pad_head(v, d);
  at::Tensor out = make_out(out_, q, q_p, d, d_og);
  at::Tensor lse = torch::empty({b, h, sq}, q.options().dtype(torch::kFloat32));
  c10::cuda::CUDAGuard guard(q.device());
  if (sk > 0 && sq > 0) {
    FaFwdParams P = {};
    P.q = q_p.data_ptr(); P.k = k_p.data_ptr(); P.v = v_p.data_ptr(); P.o = out.data_ptr(); P.softmax_lse = lse.data_ptr<float>();
    P.q_batch_stride = q_p.stride(0); P.q_row_stride = q_p.stride(1); P.q_head_stride = q_p.stride(2);
    P.k_batch_stride = k_p.stride(0); P.k_row_stride = k_p.stride(1); P.k_head_stride = k_p.stride(2);
    P.v_batch_stride = v_p.stride(0); P.v_row_stride = v_p.stride(1); P.v_head_stride = v_p.stride(2);
    P.o_batch_stride = out.stride(0); P.o_row_stride = out.stride(1); P.o_head_stride = out.stride(2);
    P.b = (int)b; P.seqlen_q = (int)sq; P.seqlen_k = (int)sk; P.h = (int)h; P.h_k = (int)hk; P.d = (int)d;
    P.is_bf16 = q.dtype() == torch::kBFloat16; P.softmax_scale = (float)softmax_scale;
    P.is_causal = is_causal; P.window_left = (int)window_size_left; P.window_right = (int)window_size_right;
    P.is_seqlens_k_cumulative = 1;
    set_alibi(P, alibi_slopes_, b, h, q.device());
    run(P, /*op=*/0, fa_b200_fwd, q);
  } else {
    out.zero_(); lse.fill_(std::numeric_limits<float>::infinity());          // flash_api.cpp:401-405
  }
  return finish(out, out_, q_p, k_p, v_p, lse, q, d, d_og);
}

// mha_varlen_fwd (flash_api.cpp:422-669).  q (total_q,h,d); k,v (total_k,hk,d) or, with block_table (b,max_blocks),
// paged (num_blocks,page,hk,d).  softmax_lse is (b,h,max_seqlen_q).
std::vector<at::Tensor> varlen_fwd(at::Tensor& q, const at::Tensor& k, const at::Tensor& v, c10::optional<at::Tensor>& out_,
                                   const at::Tensor& cu_seqlens_q, const at::Tensor& cu_seqlens_k, c10::optional<at::Tensor>& seqused_k,
                                   c10::optional<at::Tensor>& block_table_, c10::optional<at::Tensor>& alibi_slopes_,
                                   int64_t max_seqlen_q, const int64_t max_seqlen_k, const double p_dropout, const double softmax_scale,
                                   const bool zero_tensors, bool is_causal, int64_t window_size_left, int64_t window_size_right,
                                   const bool return_softmax, c10::optional<at::Generator> gen_) {
  check_qkv(q, k, v);
  TORCH_CHECK(p_dropout == 0.0, "This flash attention build does not support dropout.");
  const at::Device dev = q.device();
  check_i32(cu_seqlens_q, "cu_seqlens_q", dev); check_i32(cu_seqlens_k, "cu_seqlens_k", dev);
  TORCH_CHECK(cu_seqlens_q.is_contiguous() && cu_seqlens_k.is_contiguous(), "cu_seqlens must be contiguous");
  const bool paged = block_table_.has_value();
  TORCH_CHECK(q.dim() == 3, "q must be (total_q, nheads, headdim)");
  const int64_t total_q = q.size(0), h = q.size(1), d_og = q.size(2), b = cu_seqlens_q.numel() - 1;
  TORCH_CHECK(b > 0, "batch size must be positive");
  TORCH_CHECK(cu_seqlens_k.numel() == b + 1, "cu_seqlens_k must have shape (batch_size + 1)");
  TORCH_CHECK(d_og <= 256, "FlashAttention forward only supports head dimension at most 256");
  int64_t hk = 0, num_blocks = 0, page = 0;
  if (paged) {
    check_i32(*block_table_, "block_table", dev);
    TORCH_CHECK(block_table_->stride(-1) == 1, "block_table must have contiguous last dimension");
    TORCH_CHECK(k.dim() == 4 && v.dim() == 4, "paged k/v must be (num_blocks, page_block_size, nheads_k, headdim)");
    num_blocks = k.size(0); page = k.size(1); hk = k.size(2);
    TORCH_CHECK(page % 16 == 0, "Paged KV cache block size must be divisible by 16");
    TORCH_CHECK(block_table_->dim() == 2 && block_table_->size(0) == b, "block_table must have shape (batch_size, max_num_blocks_per_seq)");
    TORCH_CHECK(v.sizes() == k.sizes() && k.size(3) == d_og, "k/v shape mismatch");
  } else {
    TORCH_CHECK(k.dim() == 3 && v.dim() == 3, "k, v must be (total_k, nheads_k, headdim)");
    hk = k.size(1);
    TORCH_CHECK(v.sizes() == k.sizes() && k.size(2) == d_og, "k/v shape mismatch");
  }
  TORCH_CHECK(h % hk == 0, "Number of heads in key/value must divide number of heads in query");
  if (seqused_k.has_value()) {
    check_i32(*seqused_k, "seqused_k", dev);
    TORCH_CHECK(seqused_k->is_contiguous() && seqused_k->numel() == b, "seqused_k must have shape (batch_size)");
  }
  const int64_t d = (d_og + 7) / 8 * 8;
  at::Tensor q_p = pad_head(q, d), k_p = pad_head(k, d), v_p = pad_head(v, d);
  at::Tensor out = make_out(out_, q, q_p, d, d_og);
  at::Tensor lse = torch::empty({b, h, max_seqlen_q}, q.options().dtype(torch::kFloat32));
  if (zero_tensors) { out.zero_(); lse.fill_(-std::numeric_limits<float>::infinity()); }
  c10::cuda::CUDAGuard guard(dev);
  if (max_seqlen_k > 0 && max_seqlen_q > 0 && total_q > 0) {
    FaFwdParams P = {};
    P.q = q_p.data_ptr(); P.k = k_p.data_ptr(); P.v = v_p.data_ptr(); P.o = out.data_ptr(); P.softmax_lse = lse.data_ptr<float>();
    P.q_row_stride = q_p.stride(0); P.q_head_stride = q_p.stride(1);
    P.o_row_stride = out.stride(0); P.o_head_stride = out.stride(1);
    if (paged) {
      P.k_batch_stride = k_p.stride(0); P.k_row_stride = k_p.stride(1); P.k_head_stride = k_p.stride(2);
      P.v_batch_stride = v_p.stride(0); P.v_row_stride = v_p.stride(1); P.v_head_stride = v_p.stride(2);
      P.block_table = block_table_->data_ptr<int32_t>(); P.block_table_batch_stride = block_table_->stride(0);
      P.page_block_size = (int)page; P.max_blocks_per_seq = (int)block_table_->size(1); P.num_blocks = (int)num_blocks;
      P.total_k = 0;
    } else {
      P.k_row_stride = k_p.stride(0); P.k_head_stride = k_p.stride(1);
      P.v_row_stride = v_p.stride(0); P.v_head_stride = v_p.stride(1);
      P.total_k = k_p.size(0);
    }
    P.total_q = total_q;
    P.b = (int)b; P.seqlen_q = (int)max_seqlen_q; P.seqlen_k = (int)max_seqlen_k; P.h = (int)h; P.h_k = (int)hk; P.d = (int)d;
    P.is_bf16 = q.dtype() == torch::kBFloat16; P.softmax_scale = (float)softmax_scale;
    P.is_causal = is_causal; P.window_left = (int)window_size_left; P.window_right = (int)window_size_right;
    P.cu_seqlens_q = cu_seqlens_q.data_ptr<int32_t>(); P.cu_seqlens_k = cu_seqlens_k.data_ptr<int32_t>();
    P.seqused_k = seqused_k.has_value() ? seqused_k->data_ptr<int32_t>() : nullptr;
    P.is_seqlens_k_cumulative = 1;
    set_alibi(P, alibi_slopes_, b, h, dev);
    run(P, /*op=*/1, fa_b200_varlen_fwd, q);
  } else {
    out.zero_(); lse.fill_(std::numeric_limits<float>::infinity());
  }
  return finish(out, out_, q_p, k_p, v_p, lse, q, d, d_og);
}

// mha_fwd_kvcache (flash_api.cpp:671-937).  q (b,sq,h,d); cache (b_c,sk,hk,d) or paged (num_blocks,page,hk,d) with
// block_table (b,max_blocks); optional new k/v (b,sk_new,hk,d) appended in place at cache_seqlens (with rotary on k
// and q).  Returns {out, softmax_lse}.
std::vector<at::Tensor> fwd_kvcache(at::Tensor& q, const at::Tensor& kcache, const at::Tensor& vcache,
                                    c10::optional<const at::Tensor>& k_, c10::optional<const at::Tensor>& v_,
                                    c10::optional<const at::Tensor>& seqlens_k_, c10::optional<const at::Tensor>& rotary_cos_,
                                    c10::optional<const at::Tensor>& rotary_sin_, c10::optional<const at::Tensor>& cache_batch_idx_,
                                    c10::optional<at::Tensor>& block_table_, c10::optional<at::Tensor>& alibi_slopes_,
                                    c10::optional<at::Tensor>& out_, const double softmax_scale, bool is_causal,
                                    int64_t window_size_left, int64_t window_size_right, bool is_rotary_interleaved, int64_t num_splits) {
  check_qkv(q, kcache, vcache);
  const at::Device dev = q.device();
  TORCH_CHECK(q.dim() == 4 && kcache.dim() == 4 && vcache.dim() == 4, "q, kcache, vcache must be 4-D");
  const int64_t b = q.size(0), sq = q.size(1), h = q.size(2), d = q.size(3);
  const bool paged = block_table_.has_value();
  int64_t num_blocks = 0, page = 0, max_blocks = 0, hk = kcache.size(2), sk = 0, b_c = 0;
  if (paged) {
    TORCH_CHECK(!cache_batch_idx_.has_value(), "Paged KVcache does not support cache_batch_idx");
    check_i32(*block_table_, "block_table", dev);
    TORCH_CHECK(block_table_->stride(-1) == 1, "block_table must have contiguous last dimension");
    num_blocks = kcache.size(0); page = kcache.size(1);
    TORCH_CHECK(page % 16 == 0, "Paged KV cache block size must be divisible by 16");
    TORCH_CHECK(block_table_->dim() == 2 && block_table_->size(0) == b, "block_table must have shape (batch_size, max_num_blocks_per_seq)");
    max_blocks = block_table_->size(1);
    sk = max_blocks * page;
  } else {
    b_c = kcache.size(0); sk = kcache.size(1);
    if (!cache_batch_idx_.has_value()) TORCH_CHECK(b_c == b, "kcache batch size must match q (or pass cache_batch_idx)");
  }
  TORCH_CHECK(b > 0, "batch size must be postive");
  TORCH_CHECK(d <= 256, "FlashAttention forward only supports head dimension at most 256");
  TORCH_CHECK(h % hk == 0, "Number of heads in key/value must divide number of heads in query");
  TORCH_CHECK(vcache.sizes() == kcache.sizes() && kcache.size(3) == d, "kcache/vcache shape mismatch");
  TORCH_CHECK(d % 8 == 0, "fwd_kvcache: head dimension must be a multiple of 8 (cache cannot be padded in place)");
  at::Tensor out;
  if (out_.has_value()) {
    TORCH_CHECK(out_->dtype() == q.dtype(), "Output must have the same dtype as inputs");
    TORCH_CHECK(out_->is_cuda() && out_->device() == dev, "Output must be on the same device");
    TORCH_CHECK(out_->stride(-1) == 1, "Output tensor must have contiguous last dimension");
    // the reference checks `out` against its GQA-swapped shape (flash_api.cpp:787); the natural shape is accepted here
    TORCH_CHECK(out_->sizes() == q.sizes(), "out has the wrong shape");
    out = *out_;
  } else {
    out = torch::empty_like(q);
  }
  at::Tensor lse = torch::empty({b, h, sq}, q.options().dtype(torch::kFloat32));
  if (seqlens_k_.has_value()) {
    check_i32(*seqlens_k_, "seqlens_k", dev);
    TORCH_CHECK(seqlens_k_->is_contiguous() && seqlens_k_->dim() == 1 && seqlens_k_->size(0) == b, "seqlens_k must have shape (batch_size)");
  }
  if (cache_batch_idx_.has_value()) {
    check_i32(*cache_batch_idx_, "cache_batch_idx", dev);
    TORCH_CHECK(cache_batch_idx_->is_contiguous() && cache_batch_idx_->dim() == 1 && cache_batch_idx_->size(0) == b, "cache_batch_idx must have shape (batch_size)");
  }
  c10::cuda::CUDAGuard guard(dev);
  FaFwdParams P = {};
  P.q = q.data_ptr(); P.k = kcache.data_ptr(); P.v = vcache.data_ptr(); P.o = out.data_ptr(); P.softmax_lse = lse.data_ptr<float>();
  P.q_batch_stride = q.stride(0); P.q_row_stride = q.stride(1); P.q_head_stride = q.stride(2);
  P.k_batch_stride = kcache.stride(0); P.k_row_stride = kcache.stride(1); P.k_head_stride = kcache.stride(2);
  P.v_batch_stride = vcache.stride(0); P.v_row_stride = vcache.stride(1); P.v_head_stride = vcache.stride(2);
  P.o_batch_stride = out.stride(0); P.o_row_stride = out.stride(1); P.o_head_stride = out.stride(2);
  P.b = (int)b; P.seqlen_q = (int)sq; P.seqlen_k = (int)sk; P.h = (int)h; P.h_k = (int)hk; P.d = (int)d;
  P.is_bf16 = q.dtype() == torch::kBFloat16; P.softmax_scale = (float)softmax_scale;
  P.is_causal = is_causal; P.window_left = (int)window_size_left; P.window_right = (int)window_size_right;
  P.cu_seqlens_k = seqlens_k_.has_value() ? seqlens_k_->data_ptr<int32_t>() : nullptr;
  P.is_seqlens_k_cumulative = seqlens_k_.has_value() ? 0 : 1;
  P.cache_batch_idx = cache_batch_idx_.has_value() ? cache_batch_idx_->data_ptr<int32_t>() : nullptr;
  if (paged) {
    P.block_table = block_table_->data_ptr<int32_t>(); P.block_table_batch_stride = block_table_->stride(0);
    P.page_block_size = (int)page; P.max_blocks_per_seq = (int)max_blocks; P.num_blocks = (int)num_blocks;
  } else {
    P.num_blocks = (int)b_c;          // batch entries of the dense cache (extent of the K/V tensor maps)
  }
  set_alibi(P, alibi_slopes_, b, h, dev);
  P.num_splits = (int)num_splits;

  at::Tensor q_rot;
  if (k_.has_value()) {
    // ---- append new K/V into the cache (+ rotary on k and q), flash_fwd_kernel.h:672-824 ----
    TORCH_CHECK(v_.has_value(), "key and value must be passed together");
    TORCH_CHECK(seqlens_k_.has_value(), "seqlens_k must be passed if key/value are passed");
    const at::Tensor& kn = *k_; const at::Tensor& vn = *v_;
    TORCH_CHECK(kn.dtype() == q.dtype() && vn.dtype() == q.dtype(), "Key and value must have the same dtype as query");
    TORCH_CHECK(kn.is_cuda() && vn.is_cuda() && kn.stride(-1) == 1 && vn.stride(-1) == 1, "new k/v must be CUDA, last dim contiguous");
    const int64_t sk_new = kn.size(1);
    TORCH_CHECK(kn.dim() == 4 && kn.size(0) == b && kn.size(2) == hk && kn.size(3) == d && vn.sizes() == kn.sizes(), "new k/v shape mismatch");
    P.knew = kn.data_ptr(); P.vnew = vn.data_ptr(); P.seqlen_knew = (int)sk_new;
    P.knew_batch_stride = kn.stride(0); P.knew_row_stride = kn.stride(1); P.knew_head_stride = kn.stride(2);
    P.vnew_batch_stride = vn.stride(0); P.vnew_row_stride = vn.stride(1); P.vnew_head_stride = vn.stride(2);
    if (rotary_cos_.has_value()) {
      TORCH_CHECK(rotary_sin_.has_value(), "rotary_cos and rotary_sin must be passed together");
      const at::Tensor& rc = *rotary_cos_; const at::Tensor& rs = *rotary_sin_;
      TORCH_CHECK(rc.dtype() == q.dtype() && rs.dtype() == q.dtype(), "rotary_cos/sin must have the same dtype as query");
      TORCH_CHECK(rc.is_contiguous() && rs.is_contiguous(), "rotary_cos/sin must be contiguous");
      const int64_t rd = rc.size(1) * 2;
      TORCH_CHECK(rd <= d && rd % 16 == 0, "rotary_dim must be <= headdim and a multiple of 16");
      TORCH_CHECK(rc.size(0) >= sk, "cos/sin seqlen must be at least the seqlen of KV cache");
      TORCH_CHECK(rs.sizes() == rc.sizes(), "rotary_sin shape mismatch");
      P.rotary_cos = rc.data_ptr(); P.rotary_sin = rs.data_ptr();
      P.rotary_dim = (int)rd; P.seqlen_ro = (int)rc.size(0); P.is_rotary_interleaved = is_rotary_interleaved;
    }
    const int fused = fa_b200_fused_append_supported(&P);
    TORCH_CHECK(fused >= 0, fa_b200_last_error());
    if (!fused) {
      // not a decode shape (or a sliding window): the stand-alone append kernel first, then attention over the grown
      // cache with the rotated q.  Otherwise knew / vnew / rotary stay in the block and the decode kernel does it all.
      if (rotary_cos_.has_value()) {
        q_rot = torch::empty({b, sq, h, d}, q.options());
        P.workspace = q_rot.data_ptr(); P.workspace_bytes = (size_t)q_rot.numel() * 2;
      }
      ok(fa_b200_append_kv(&P, at::cuda::getCurrentCUDAStream(q.get_device()).stream()));
      P.knew = P.vnew = P.rotary_cos = P.rotary_sin = nullptr;
      P.workspace = nullptr; P.workspace_bytes = 0;
      if (q_rot.defined()) {
        P.q = q_rot.data_ptr();
        P.q_batch_stride = q_rot.stride(0); P.q_row_stride = q_rot.stride(1); P.q_head_stride = q_rot.stride(2);
      }
    }
  } else {
    TORCH_CHECK(!rotary_cos_.has_value(), "rotary embedding requires new key/value (reference applies it in the append path)");
  }
  run(P, /*op=*/2, fa_b200_fwd_kvcache, q);
  return {out, lse};
}

py::tuple last_call_info() { return py::make_tuple(g_info.kernel_used, g_info.num_splits, g_info.launches); }

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.doc() = "FlashAttention forward for B200 (libfa_b200.so) behind the reference's extension interface";
  m.def("fwd", &fwd, "Forward pass");
  m.def("varlen_fwd", &varlen_fwd, "Forward pass (variable length)");
  m.def("fwd_kvcache", &fwd_kvcache, "Forward pass, with KV-cache");
  m.def("last_call_info", &last_call_info, "(kernel_used, num_splits, launches) of the last call on this host thread");
  m.def("abi_version", []() { return fa_b200_abi_version(); });
}
