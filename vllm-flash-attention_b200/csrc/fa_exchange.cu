// fa_exchange.cu — cross-GPU merge of sequence-sharded decode partials in ONE kernel per rank:
// push this rank's (out fp32 | lse) record into every peer's receive area with plain stores over
// NVLink peer mappings, publish a sequence number, wait for the peers' numbers, LSE-combine.
// Replaces "NCCL all-gather + combine launch" (two launches and a collective's latency) for the
// long-context decode configuration; the algebra is combine_attn_seqk_parallel's
// (reference src/flash_fwd_kernel.h:1151-1221) lifted from splits to ranks.
//
// Memory (all in a symmetric allocation, same layout on every rank; see vllm_flash_attn/sharded.py):
//   recv[2][world][record]        receive area, double-buffered by call parity
//   flags[world][FA_XCHG_CTAS]    flags[r][c] = sequence number of the last slice CTA c of rank r delivered here
//   state[FA_XCHG_CTAS]           state[c] = calls CTA c of this rank has completed
// The rows of the record are dealt to the CTAs the same way on every rank, and CTA c both pushes and merges
// exactly its own rows, so it depends on the peers' CTA c only.  Protocol of CTA c, call n = state[c] + 1:
//   1. copy its rows (O row + lse) of the local record to recv[n&1][rank] of every rank (incl. self), fence
//      system-wide, store n into flags[rank][c] of every rank;
//   2. wait until flags[r][c] >= n for all r (one polling thread per peer), merge its rows from recv[n&1][*].
// No count over the CTAs of a rank, no second fence, no CTA waits for data it does not merge.  All FA_XCHG_CTAS
// CTAs run in every call (idle ones only exchange flags), so their call numbers stay in lockstep whatever
// sequence of shapes is merged.  A rank cannot run two calls ahead of a peer (its call n+1 needs the peer's flag n+1, sent only after the peer finished
// call n), so two receive buffers suffice.  All state lives in device memory: the launch is CUDA-graph
// replayable.  The wait is bounded (trap after FA_XCHG_TIMEOUT_NS = 120 s; all ranks should be warmed up before the first merge) so that a missing peer surfaces as a launch failure,
// not a hung GPU.
#include "fa_common.cuh"
#include "fa_launch.h"

namespace fa {

struct ExchangeParams {
  float* recv[FA_MAX_PEERS];
  unsigned* flags[FA_MAX_PEERS];
  const float* local;       // this rank's record [o (rows*d) | lse (rows)]
  unsigned* state;
  int rank, world;
  int b, h, sq, d;
  int64_t record;           // fp32 elements per record slot (>= rows*(d+1), multiple of 4)
  void* o; float* lse;      // outputs: o (b,sq,h,d) by strides, lse (b,h,sq) or NULL
  int64_t o_batch_stride, o_row_stride, o_head_stride;
  int out_dtype;            // 0 fp32, 1 bf16, 2 fp16
};

__global__ void __launch_bounds__(256) fa_exchange_combine_kernel(const __grid_constant__ ExchangeParams p) {
  __shared__ unsigned s_n;
  const int cta = blockIdx.x;
  if (threadIdx.x == 0) s_n = p.state[cta] + 1u;               // only this CTA ever writes state[cta]
  __syncthreads();
  const unsigned n = s_n;
  const int64_t rows = (int64_t)p.b * p.h * p.sq;
  const int64_t slot = ((int64_t)(n & 1u) * p.world + p.rank) * p.record;
  const int lane = threadIdx.x & 31, warps = blockDim.x >> 5;

  // ---- 1. push this CTA's rows: one warp per row, 16-byte pieces ----
  for (int64_t row = (int64_t)cta * warps + (threadIdx.x >> 5); row < rows; row += (int64_t)gridDim.x * warps) {
    for (int c = lane * 4; c < p.d; c += 128) {
      const float4 v = *reinterpret_cast<const float4*>(p.local + row * p.d + c);
      for (int r = 0; r < p.world; r++) *reinterpret_cast<float4*>(p.recv[r] + slot + row * p.d + c) = v;
    }
    if (lane == 0) {
      const float l = p.local[rows * p.d + row];
      for (int r = 0; r < p.world; r++) p.recv[r][slot + rows * p.d + row] = l;
    }
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x < p.world) {
    const int r = threadIdx.x;
    st_release_sys(p.flags[r] + p.rank * FA_XCHG_CTAS + cta, n);
    // ---- 2. wait for rank r's slice of call n ----
    wait_flag_sys(p.flags[p.rank] + r * FA_XCHG_CTAS + cta, n);
  }
  if (threadIdx.x == 0) p.state[cta] = n;
  __syncthreads();

  // ---- 3. merge: one warp per row ----
  const float* base = p.recv[p.rank] + (int64_t)(n & 1u) * p.world * p.record;
  for (int64_t row = (int64_t)cta * warps + (threadIdx.x >> 5); row < rows; row += (int64_t)gridDim.x * warps) {
    float ls[FA_MAX_PEERS], mx = -INFINITY;
#pragma unroll
    for (int r = 0; r < FA_MAX_PEERS; r++) {
      ls[r] = r < p.world ? __ldcg(base + r * p.record + rows * p.d + row) : -INFINITY;
      mx = fmaxf(mx, ls[r]);
    }
    const bool empty = mx == -INFINITY;
    float sum = 0.f;
#pragma unroll
    for (int r = 0; r < FA_MAX_PEERS; r++) sum += empty ? 0.f : expf(ls[r] - mx);
    const float lse = empty ? INFINITY : logf(sum) + mx;      // rows without any key: O = 0, LSE = +inf
    float w[FA_MAX_PEERS];
#pragma unroll
    for (int r = 0; r < FA_MAX_PEERS; r++) w[r] = empty ? 0.f : expf(ls[r] - lse);
    const int qi = (int)(row % p.sq), hi = (int)((row / p.sq) % p.h), bi = (int)(row / ((int64_t)p.sq * p.h));
    const int64_t o_off = (int64_t)bi * p.o_batch_stride + (int64_t)qi * p.o_row_stride + (int64_t)hi * p.o_head_stride;
    for (int c = lane * 4; c < p.d; c += 128) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int r = 0; r < FA_MAX_PEERS; r++) {
        if (r < p.world) {
          const float4 v = __ldcg(reinterpret_cast<const float4*>(base + r * p.record + row * p.d + c));
          acc.x += w[r] * v.x; acc.y += w[r] * v.y; acc.z += w[r] * v.z; acc.w += w[r] * v.w;
        }
      }
      if (p.out_dtype == 0) {
        *reinterpret_cast<float4*>(reinterpret_cast<float*>(p.o) + o_off + c) = acc;
      } else if (p.out_dtype == 1) {
        uint2 pk; pk.x = Pack2<__nv_bfloat16>::pack(acc.x, acc.y); pk.y = Pack2<__nv_bfloat16>::pack(acc.z, acc.w);
        *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(p.o) + o_off + c) = pk;
      } else {
        uint2 pk; pk.x = Pack2<__half>::pack(acc.x, acc.y); pk.y = Pack2<__half>::pack(acc.z, acc.w);
        *reinterpret_cast<uint2*>(reinterpret_cast<__half*>(p.o) + o_off + c) = pk;
      }
    }
    if (p.lse && lane == 0) p.lse[row] = lse;
  }
}

cudaError_t launch_exchange_combine(const FaExchangeParams& a, cudaStream_t stream) {
  ExchangeParams p;
  for (int r = 0; r < FA_MAX_PEERS; r++) {
    p.recv[r] = r < a.world ? reinterpret_cast<float*>(a.recv[r]) : nullptr;
    p.flags[r] = r < a.world ? a.flags[r] : nullptr;
  }
  p.local = reinterpret_cast<const float*>(a.local_record); p.state = a.state;
  p.rank = a.rank; p.world = a.world; p.b = a.b; p.h = a.h; p.sq = a.seqlen_q; p.d = a.d; p.record = a.record_elems;
  p.o = a.o; p.lse = a.softmax_lse;
  p.o_batch_stride = a.o_batch_stride; p.o_row_stride = a.o_row_stride; p.o_head_stride = a.o_head_stride;
  p.out_dtype = a.out_dtype;
  const int64_t rows = (int64_t)a.b * a.h * a.seqlen_q;
  // Always FA_XCHG_CTAS CTAs, whatever the shape: the protocol state (call number, receive-buffer parity) is kept
  // per CTA, so every CTA must take part in every call — a CTA without rows still publishes and waits for its
  // flags.  With a shape-dependent grid the CTAs' call numbers drift apart when the batch size changes between
  // calls (continuous batching) and a fast rank's call n+1 could overwrite the buffer half a slow peer is still
  // merging for call n.  Every CTA has to be resident while it waits for the peers: 64 CTAs of 256 threads are
  // far below one wave.
  (void)rows;
  fa_exchange_combine_kernel<<<FA_XCHG_CTAS, 256, 0, stream>>>(p);
  return cudaGetLastError();
}

}  // namespace fa
