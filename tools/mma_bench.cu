// tcgen05.mma rate microbenchmark (development tool): one thread issues the prefill kernel's MMA shapes back to
// back on garbage shared memory and reports SM clocks per (QK^T + PV) group.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -Ivllm-flash-attention_b200/csrc \
//        -Iinclude -o tools/mma_bench.bin tools/mma_bench.cu
#include <cstdio>
#include "fa_common.cuh"
#include "fa_sm100.cuh"
using namespace fa;
using namespace fa::sm100;

// mode 0: d=128 (8 SS N=128 + 8 TS N=128)   mode 1: d=256 (16 SS N=128 + 8 TS N=256)
// mode 2: only the 16 SS MMAs of d=256      mode 3: only the 8 TS N=256 MMAs      mode 4: 16 TS (A=Q in TMEM) N=128
__global__ void __launch_bounds__(128, 1) mma_bench(long long* clk, int mode, int iters) {
  extern __shared__ uint8_t dsm[];
  __shared__ uint32_t slot;
  __shared__ __align__(8) uint64_t bar;
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); fence_barrier_init(); }
  if (warp == 0) { tmem_alloc(smem_u32(&slot), 512); tmem_relinquish(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = slot;
  const uint32_t base = (smem_u32(dsm) + 1023u) & ~1023u;
  constexpr uint64_t DK = ((uint64_t)(16 >> 4) << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
  constexpr uint64_t DV = ((uint64_t)(16384 >> 4) << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
  if (warp == 1 && elect_one()) {
    const uint32_t iqk = make_idesc(true, 128, 128, false);
    const uint32_t ipv = make_idesc(true, 128, mode == 0 ? 128 : 256, true);
    const int nqk = mode == 0 ? 8 : 16;
    const uint32_t qa = (base & 0x3FFFF) >> 4, kb = ((base + 65536) & 0x3FFFF) >> 4, vb = ((base + 131072) & 0x3FFFF) >> 4;
    uint32_t ph = 0;
    const long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
      if (mode != 3) {
        for (int ks = 0; ks < nqk; ks++) {
          const uint32_t off = (ks >> 2) * (16384 >> 4) + (ks & 3) * 2;
          if (mode == 4) umma_ts(tmem, tmem + 128 + ks * 8, DK | (kb + off), iqk, ks > 0);
          else umma_ss(tmem, DK | (qa + off), DK | (kb + off), iqk, ks > 0);
        }
      }
      if (mode == 0 || mode == 1 || mode == 3)
        for (int kk = 0; kk < 8; kk++) umma_ts(tmem + 256, tmem + kk * 8, DV | (vb + kk * (2048 >> 4)), ipv, 1);
      umma_commit(smem_u32(&bar));
      mbar_wait(smem_u32(&bar), ph); ph ^= 1;
    }
    clk[0] = clock64() - t0;
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

int main() {
  long long* clk; cudaMalloc(&clk, 8);
  cudaFuncSetAttribute(mma_bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const char* names[] = {"d=128: 8 QK (SS N=128) + 8 PV (TS N=128)", "d=256: 16 QK (SS N=128) + 8 PV (TS N=256)", "16 QK SS N=128 only",
                         "8 PV TS N=256 only", "16 QK with A in TMEM (TS N=128)"};
  for (int mode = 0; mode < 5; mode++) {
    for (int rep = 0; rep < 2; rep++) mma_bench<<<1, 128, 200 * 1024>>>(clk, mode, 500);
    long long h = 0; cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);
    printf("%-46s %.0f clk per group\n", names[mode], (double)h / 500);
  }
  printf("status: %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
}
