// Instruction-throughput microbenchmark for the softmax inner loop (development tool).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/microbench tools/microbench.cu && /tmp/microbench
// One CTA, W warps per SMSP; each thread runs N independent chains of one operation for ITER
// iterations; prints SM clocks per warp-instruction per SMSP.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int ITER = 2048;
constexpr int N = 8;   // independent chains per thread

__device__ __forceinline__ float ex2(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

template <int OP>
__global__ void bench(float* out, long long* clk, float seed) {
  float a[N], b[N];
  float2 p[N], q[N];
#pragma unroll
  for (int i = 0; i < N; i++) { a[i] = seed + i + threadIdx.x * 1e-3f; b[i] = seed * 0.5f + i; p[i] = make_float2(a[i], b[i]); q[i] = make_float2(b[i], a[i]); }
  __syncthreads();
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITER; it++) {
#pragma unroll
    for (int i = 0; i < N; i++) {
      if (OP == 0) a[i] = fmaf(a[i], b[i], 1.0001f * b[i]);                         // FFMA 3-reg
      if (OP == 1) p[i] = __ffma2_rn(p[i], q[i], q[i]);                              // FFMA2
      if (OP == 2) p[i] = __fadd2_rn(p[i], q[i]);                                    // FADD2
      if (OP == 3) a[i] = ex2(a[i]);                                                 // MUFU.EX2
      if (OP == 4) { uint32_t r; asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(a[i]), "f"(b[i])); a[i] = __uint_as_float(r); }  // F2FP
      if (OP == 5) a[i] = fmaxf(a[i], fmaxf(b[i], p[i].x));                          // FMNMX3
      if (OP == 6) a[i] = __int_as_float(__float_as_int(a[i]) + (__float_as_int(b[i]) << 23));  // exponent add
      if (OP == 7) a[i] = a[i] + b[i];                                               // FADD
      if (OP == 8) a[i] = fmaxf(a[i], b[i]);                                         // FMNMX
      if (OP == 9) { a[i] = ex2(a[i]); p[i] = __ffma2_rn(p[i], q[i], q[i]); }        // MUFU + FFMA2 co-issue
    }
  }
  long long t1 = clock64();
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < N; i++) s += a[i] + b[i] + p[i].x + p[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) clk[0] = t1 - t0;
}

template <int OP> void run(const char* name, int ops_per_iter_elem) {
  float* out; long long* clk;
  cudaMalloc(&out, 1024 * 4); cudaMalloc(&clk, 8);
  for (int warps_per_smsp = 1; warps_per_smsp <= 2; warps_per_smsp++) {
    bench<OP><<<1, 128 * warps_per_smsp>>>(out, clk, 0.37f);
    bench<OP><<<1, 128 * warps_per_smsp>>>(out, clk, 0.37f);
    long long h; cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);
    double per = (double)h / (ITER * N * ops_per_iter_elem * warps_per_smsp);
    printf("%-28s warps/SMSP=%d  %.2f clk per warp-instruction per SMSP\n", name, warps_per_smsp, per);
  }
  cudaFree(out); cudaFree(clk);
}

int main() {
  run<0>("FFMA (3 reg)", 1);
  run<1>("FFMA2 (packed f32x2)", 1);
  run<2>("FADD2 (packed f32x2)", 1);
  run<3>("MUFU.EX2", 1);
  run<4>("F2FP.BF16 pack", 1);
  run<5>("FMNMX3", 1);
  run<6>("SHL+IADD exponent add", 1);
  run<7>("FADD", 1);
  run<8>("FMNMX", 1);
  run<9>("MUFU.EX2 + FFMA2 pair", 1);
  cudaError_t e = cudaDeviceSynchronize();
  printf("status: %s\n", cudaGetErrorString(e));
  return 0;
}
