// Softmax-warp microbenchmark (development tool): the per-tile instruction sequence of the tcgen05 prefill
// kernel's softmax warps (LDTM 128 fp32 scores -> FFMA2 -> EX2 -> row sum + bf16 pack -> STTM P), with no
// MMA and no barriers, to find the instruction schedule a lone warp per sub-partition runs fastest.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -Ivllm-flash-attention_b200/csrc \
//        -Iinclude -DVARIANT=0 -o tools/softmax_bench.bin tools/softmax_bench.cu
// Prints SM clocks per 128-column tile for 1 and 2 warps per sub-partition.
#include <cstdio>
#include "fa_common.cuh"
#include "fa_sm100.cuh"

using namespace fa;
using namespace fa::sm100;

#ifndef VARIANT
#define VARIANT 0
#endif
constexpr int TILES = 2000;

__device__ __forceinline__ void st16(uint32_t addr, const uint32_t (&pk)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
      ::"r"(addr), "r"(pk[0]), "r"(pk[1]), "r"(pk[2]), "r"(pk[3]), "r"(pk[4]), "r"(pk[5]), "r"(pk[6]),
        "r"(pk[7]), "r"(pk[8]), "r"(pk[9]), "r"(pk[10]), "r"(pk[11]), "r"(pk[12]), "r"(pk[13]), "r"(pk[14]), "r"(pk[15])
      : "memory");
}

template <typename T>
__global__ void __launch_bounds__(384, 1) bench(float* out, long long* clk, float mul, float m_ref0, int workers, int mma_on) {
  __shared__ uint32_t slot;
  __shared__ __align__(8) uint64_t spin_bar, mma_bar;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&spin_bar), workers * 32); mbar_init(smem_u32(&mma_bar), 1); fence_barrier_init(); }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) { tmem_alloc(smem_u32(&slot), 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = slot;
  const int t = warp >> 2, wq = warp & 3;
  const uint32_t tm_s = tmem + ((uint32_t)(wq * 32) << 16) + t * 128;
  // fill S with moderate scores
  if (warp < 8) {
    uint32_t r[32];
#pragma unroll
    for (int c = 0; c < 4; c++) {
#pragma unroll
      for (int i = 0; i < 32; i++) r[i] = __float_as_uint(-0.01f * (float)((i * 7 + c * 3 + lane) & 63));
      tmem_st32(tm_s + 256 + c * 32, r);     // keep a pristine copy in the O region: P overwrites S
    }
    tmem_st_wait();
  }
  __syncthreads();
  if (warp >= workers) {      // spinner warps: what the producer / issuer / idle softmax warps do in the real kernel
    if (warp == workers && mma_on) {
      // one thread keeps the tensor pipe busy with the real kernel's MMA shapes: QK-like SS MMAs into S1 and PV-like
      // TS MMAs (A from TMEM) into O1, on whatever shared memory holds
      extern __shared__ uint8_t dsm[];
      const uint32_t base = (smem_u32(dsm) + 1023u) & ~1023u;
      constexpr uint64_t DK = ((uint64_t)(16 >> 4) << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
      constexpr uint64_t DV = ((uint64_t)(16384 >> 4) << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
      constexpr uint32_t iqk = make_idesc(true, 128, 128, false), ipv = make_idesc(true, 128, 128, true);
      if (elect_one()) {
        uint32_t ph = 0;
        while (!mbar_try_wait(smem_u32(&spin_bar), 0)) {
          const uint32_t qa = (base & 0x3FFFF) >> 4, kb = ((base + 32768) & 0x3FFFF) >> 4, vb = ((base + 65536) & 0x3FFFF) >> 4;
#pragma unroll
          for (int ks = 0; ks < 8; ks++) {
            const uint32_t off = (ks >> 2) * (16384 >> 4) + (ks & 3) * 2;
            umma_ss(tmem + 128, DK | (qa + off), DK | (kb + off), iqk, ks > 0);
          }
#pragma unroll
          for (int kk = 0; kk < 8; kk++) umma_ts(tmem + 384, tmem + 128 + kk * 8, DV | (vb + kk * (2048 >> 4)), ipv, 1);
          umma_commit(smem_u32(&mma_bar));
          mbar_wait(smem_u32(&mma_bar), ph); ph ^= 1;
        }
      }
      __syncwarp();
    } else {
      if (lane == 0) mbar_wait(smem_u32(&spin_bar), 0);
      __syncwarp();
    }
  } else {
  float m_ref = m_ref0, l_sum = 0.f;
  const float2 mul2 = make_float2(mul, mul);
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < TILES; it++) {
    uint32_t sr[4][32];
    tmem_ld32(tm_s + 256, sr[0]); tmem_ld32(tm_s + 256 + 32, sr[1]);
    tmem_ld_wait();
    tmem_ld32(tm_s + 256 + 64, sr[2]); tmem_ld32(tm_s + 256 + 96, sr[3]);
    float2 rsa = make_float2(0.f, 0.f), rsb = make_float2(0.f, 0.f);
    const float2 nms2 = make_float2(-m_ref * mul, -m_ref * mul);
    auto ex = [&](int c, float2 (&pp)[16]) {
#pragma unroll
      for (int i = 0; i < 16; i++) {
        const float2 x = __ffma2_rn(make_float2(__uint_as_float(sr[c][2 * i]), __uint_as_float(sr[c][2 * i + 1])), mul2, nms2);
        pp[i] = make_float2(fast_exp2(x.x), fast_exp2(x.y));
      }
    };
    auto pack = [&](int c, const float2 (&pp)[16]) {
      uint32_t pk[16];
#pragma unroll
      for (int i = 0; i < 16; i++) {
        if (i & 1) rsb = __fadd2_rn(rsb, pp[i]); else rsa = __fadd2_rn(rsa, pp[i]);
        pk[i] = Pack2<T>::pack(pp[i].x, pp[i].y);
      }
      st16(tm_s + c * 16, pk);
    };
#if VARIANT == 0
    {   // the kernel's current source order
      float2 pa[16], pb[16];
      ex(0, pa); ex(1, pb); pack(0, pa);
      tmem_ld_wait();
      ex(2, pa); pack(1, pb);
      ex(3, pb); pack(2, pa); pack(3, pb);
    }
#elif VARIANT == 1
    {   // same, stages separated by __syncwarp (does ptxas keep consumers a stage behind?)
      float2 pa[16], pb[16];
      ex(0, pa); __syncwarp(); ex(1, pb); pack(0, pa); __syncwarp();
      tmem_ld_wait();
      ex(2, pa); pack(1, pb); __syncwarp();
      ex(3, pb); pack(2, pa); __syncwarp(); pack(3, pb);
    }
#elif VARIANT == 2
    {   // exponentials written in place as b32 through asm volatile so the consumer order is the source order
      tmem_ld_wait();
      uint32_t e[4][32];
      auto exv = [&](int c) {
#pragma unroll
        for (int i = 0; i < 16; i++) {
          const float2 x = __ffma2_rn(make_float2(__uint_as_float(sr[c][2 * i]), __uint_as_float(sr[c][2 * i + 1])), mul2, nms2);
          asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=r"(e[c][2 * i]) : "f"(x.x));
          asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=r"(e[c][2 * i + 1]) : "f"(x.y));
        }
      };
      auto packv = [&](int c) {
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 16; i++) {
          const float2 pp = make_float2(__uint_as_float(e[c][2 * i]), __uint_as_float(e[c][2 * i + 1]));
          if (i & 1) rsb = __fadd2_rn(rsb, pp); else rsa = __fadd2_rn(rsa, pp);
          asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(pk[i]) : "f"(pp.y), "f"(pp.x));
        }
        st16(tm_s + c * 16, pk);
      };
      exv(0); exv(1); packv(0); exv(2); packv(1); exv(3); packv(2); packv(3);
    }
#elif VARIANT == 3
    {   // pair-granular software pipeline: pack of pair i-LAG interleaved with exp of pair i, all asm volatile
      tmem_ld_wait();
      constexpr int LAG = 12;
      uint32_t e[128];
      uint32_t pk[64];
#pragma unroll
      for (int i = 0; i < 64 + LAG; i++) {
        if (i < 64) {
          const int c = i >> 4, k = i & 15;
          const float2 x = __ffma2_rn(make_float2(__uint_as_float(sr[c][2 * k]), __uint_as_float(sr[c][2 * k + 1])), mul2, nms2);
          asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=r"(e[2 * i]) : "f"(x.x));
          asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=r"(e[2 * i + 1]) : "f"(x.y));
        }
        if (i >= LAG) {
          const int j = i - LAG;
          const float2 pp = make_float2(__uint_as_float(e[2 * j]), __uint_as_float(e[2 * j + 1]));
          if (j & 1) rsb = __fadd2_rn(rsb, pp); else rsa = __fadd2_rn(rsa, pp);
          asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(pk[j]) : "f"(pp.y), "f"(pp.x));
          if ((j & 15) == 15) {
            const int c = j >> 4;
            uint32_t q[16];
#pragma unroll
            for (int u = 0; u < 16; u++) q[u] = pk[c * 16 + u];
            st16(tm_s + c * 16, q);
          }
        }
      }
    }
#endif
    tmem_st_wait();
    const float rs = (rsa.x + rsa.y) + (rsb.x + rsb.y);
    l_sum += rs;
    if (__any_sync(0xffffffffu, rs > 1e30f)) m_ref += 1.f;     // never taken; keeps the check in the loop
  }
  const long long t1 = clock64();
  out[threadIdx.x] = l_sum + m_ref;
  if (threadIdx.x == 0 && blockIdx.x == 0) clk[0] = t1 - t0;
  mbar_arrive(smem_u32(&spin_bar));
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

int main() {
  float* out; long long* clk;
  cudaMalloc(&out, 4096); cudaMalloc(&clk, 8);
  cudaFuncSetAttribute(bench<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  for (int grid : {1, 148})
  for (int wps = 1; wps <= 2; wps++)
  for (int mma = 0; mma <= 1; mma++) {
    const int spin = 1;     // 12 warps as in the kernel: wps*4 working, the rest wait on an mbarrier (one of them issues MMAs)
    for (int rep = 0; rep < 2; rep++) bench<__nv_bfloat16><<<grid, 384, 100 * 1024>>>(out, clk, 0.1275f, 0.f, 4 * wps, mma);
    long long h = 0; cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);
    printf("VARIANT %d grid=%d working warps/SMSP=%d mma=%d  %.0f clk per tile (per warp), %.0f clk per tile per SMSP\n", VARIANT, grid, wps, mma,
           (double)h / TILES, (double)h / TILES / wps);
    (void)spin;
  }
  printf("status: %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
