// fa_warp_attn.cuh — one warp's online-softmax attention over KV chunks held in shared memory,
// on mma.sync m16n8k16.  Shared by the decode kernel (fa_decode.cu: warps split the KV range,
// M = packed GQA rows) and the tile kernel (fa_tile.cu: warps split the query rows).
//
// Arithmetic follows the reference's online softmax (src/softmax.h:66-186): fp32 scores,
// p = exp2(s*scale*log2e - max), row sum on the unrounded p, P rounded to the I/O dtype before
// P.V (flash_fwd_kernel.h:323,382), O rescaled by exp2(m_old - m_new).
#pragma once
#include "fa_common.cuh"

namespace fa {

// Shared-memory tile layout: `rows` rows of D elements (2 B), 16-byte chunks XOR-swizzled so that
// the 8 row addresses of an ldmatrix phase fall in 8 distinct 16 B bank groups.
template <int D> __device__ __forceinline__ uint32_t swz(int row, int chunk) {
  constexpr int CPR = D / 8;
  if constexpr (CPR % 8 == 0) return (uint32_t)(row * (D * 2) + ((chunk ^ (row & 7)) << 4));
  else                        return (uint32_t)(row * (D * 2) + ((chunk ^ ((row >> 1) & 3)) << 4));
}

template <typename T, int D, int MT, int NC>
struct WarpAttn {
  static constexpr int NT_S = NC / 8;   // score n-tiles per step
  static constexpr int NT_O = D / 8;    // output n-tiles
  float o[MT][NT_O][4];
  float m[MT][2];   // running max of scaled (log2-domain) scores; row lane/4 and lane/4+8
  float l[MT][2];   // running sum (per-thread partial over its columns; quad-reduced at the end)

  __device__ __forceinline__ void init() {
#pragma unroll
    for (int i = 0; i < MT; i++) {
      m[i][0] = m[i][1] = -INFINITY; l[i][0] = l[i][1] = 0.f;
#pragma unroll
      for (int j = 0; j < NT_O; j++) { o[i][j][0] = o[i][j][1] = o[i][j][2] = o[i][j][3] = 0.f; }
    }
  }

  // One step over NC keys.  q_s / k_s / v_s: shared addresses of swizzled tiles (Q: 16*MT rows;
  // K, V: NC rows).  mask(row, col, x) -> x' is applied to the scaled score when need_mask;
  // row in [0,16*MT), col in [0,NC); it returns -INFINITY for hidden keys (and may add ALiBi).
  template <class MaskF>
  __device__ __forceinline__ void step(uint32_t q_s, uint32_t k_s, uint32_t v_s, float scale_log2,
                                       bool need_mask, MaskF mask) {
    const int lane = threadIdx.x & 31;
    float s[MT][NT_S][4];
#pragma unroll
    for (int i = 0; i < MT; i++)
#pragma unroll
      for (int j = 0; j < NT_S; j++) s[i][j][0] = s[i][j][1] = s[i][j][2] = s[i][j][3] = 0.f;

    // ---- S = Q K^T ----
#pragma unroll
    for (int ks = 0; ks < D / 16; ks++) {
      uint32_t a[MT][4];
#pragma unroll
      for (int i = 0; i < MT; i++) {
        int row = i * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
        ldsm_x4(q_s + swz<D>(row, ks * 2 + (lane >> 4)), a[i][0], a[i][1], a[i][2], a[i][3]);
      }
#pragma unroll
      for (int jp = 0; jp < NT_S / 2; jp++) {
        uint32_t b0, b1, b2, b3;
        int row = jp * 16 + (lane & 7) + (lane >> 4) * 8;
        ldsm_x4(k_s + swz<D>(row, ks * 2 + ((lane >> 3) & 1)), b0, b1, b2, b3);
#pragma unroll
        for (int i = 0; i < MT; i++) {
          mma16816<T>(s[i][2 * jp], a[i], b0, b1);
          mma16816<T>(s[i][2 * jp + 1], a[i], b2, b3);
        }
      }
    }

    // ---- scale (+mask), online softmax ----
    uint32_t pa[MT][NT_S / 2][4];
#pragma unroll
    for (int i = 0; i < MT; i++) {
      float mx[2] = {m[i][0], m[i][1]};
#pragma unroll
      for (int j = 0; j < NT_S; j++)
#pragma unroll
        for (int e = 0; e < 4; e++) {
          float x = s[i][j][e] * scale_log2;
          if (need_mask) x = mask(i * 16 + (lane >> 2) + (e >> 1) * 8, j * 8 + (lane & 3) * 2 + (e & 1), x);
          s[i][j][e] = x;
          mx[e >> 1] = fmaxf(mx[e >> 1], x);
        }
#pragma unroll
      for (int r = 0; r < 2; r++) {
        mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
        mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
      }
      float alpha[2], msafe[2];
#pragma unroll
      for (int r = 0; r < 2; r++) {
        msafe[r] = mx[r] == -INFINITY ? 0.f : mx[r];          // fully masked so far: avoid inf-inf
        alpha[r] = fast_exp2(m[i][r] - msafe[r]);             // m = -inf -> 0
        m[i][r] = mx[r];
        l[i][r] *= alpha[r];
      }
#pragma unroll
      for (int j = 0; j < NT_S; j++) {
        float p0 = fast_exp2(s[i][j][0] - msafe[0]), p1 = fast_exp2(s[i][j][1] - msafe[0]);
        float p2 = fast_exp2(s[i][j][2] - msafe[1]), p3 = fast_exp2(s[i][j][3] - msafe[1]);
        l[i][0] += p0 + p1; l[i][1] += p2 + p3;
        pa[i][j >> 1][(j & 1) * 2 + 0] = Pack2<T>::pack(p0, p1);
        pa[i][j >> 1][(j & 1) * 2 + 1] = Pack2<T>::pack(p2, p3);
      }
#pragma unroll
      for (int j = 0; j < NT_O; j++) {
        o[i][j][0] *= alpha[0]; o[i][j][1] *= alpha[0];
        o[i][j][2] *= alpha[1]; o[i][j][3] *= alpha[1];
      }
    }

    // ---- O += P V ----
#pragma unroll
    for (int kk = 0; kk < NC / 16; kk++) {
#pragma unroll
      for (int jp = 0; jp < NT_O / 2; jp++) {
        uint32_t b0, b1, b2, b3;
        int row = kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
        ldsm_x4_t(v_s + swz<D>(row, jp * 2 + (lane >> 4)), b0, b1, b2, b3);
#pragma unroll
        for (int i = 0; i < MT; i++) {
          mma16816<T>(o[i][2 * jp], pa[i][kk], b0, b1);
          mma16816<T>(o[i][2 * jp + 1], pa[i][kk], b2, b3);
        }
      }
    }
  }

  // Quad-reduce the row sums (call once after the last step).
  __device__ __forceinline__ void finish_sums() {
#pragma unroll
    for (int i = 0; i < MT; i++)
#pragma unroll
      for (int r = 0; r < 2; r++) {
        l[i][r] += __shfl_xor_sync(0xffffffffu, l[i][r], 1);
        l[i][r] += __shfl_xor_sync(0xffffffffu, l[i][r], 2);
      }
  }
};

// Visibility window of query position `qpos` (bottom-right aligned; reference src/mask.h:171-172):
// keys [lo, hi) are visible.
struct RowWindow { int lo, hi; };
__device__ __forceinline__ RowWindow row_window(const KParams& p, int qpos, int sq, int sk) {
  RowWindow w; w.lo = 0; w.hi = sk;
  int shift = qpos + sk - sq;
  if (p.is_causal) { w.hi = min(sk, shift + 1); }
  else if (p.is_local) {
    w.hi = min(sk, shift + 1 + p.window_right);
    w.lo = max(0, shift - p.window_left);
  }
  return w;
}

}  // namespace fa
