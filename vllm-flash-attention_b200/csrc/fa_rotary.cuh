// fa_rotary.cuh — row copy with rotary embedding (reference src/rotary.h:19-145), shared by the stand-alone append kernel
// (fa_append.cu) and the decode kernel's in-kernel append (fa_decode.cu).  Rotation in fp32, rounded once.
#pragma once
#include "fa_common.cuh"

namespace fa {

template <typename T> __device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]);
template <> __device__ __forceinline__ void unpack8<__nv_bfloat16>(const uint4& u, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; i++) { f[2 * i] = __low2float(h[i]); f[2 * i + 1] = __high2float(h[i]); }
}
template <> __device__ __forceinline__ void unpack8<__half>(const uint4& u, float (&f)[8]) {
  const __half2* h = reinterpret_cast<const __half2*>(&u);
#pragma unroll
  for (int i = 0; i < 4; i++) { f[2 * i] = __low2float(h[i]); f[2 * i + 1] = __high2float(h[i]); }
}
template <typename T> __device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 u;
  u.x = Pack2<T>::pack(f[0], f[1]); u.y = Pack2<T>::pack(f[2], f[3]);
  u.z = Pack2<T>::pack(f[4], f[5]); u.w = Pack2<T>::pack(f[6], f[7]);
  return u;
}

// Copy one row of d elements src -> dst, rotating the first rotary_dim elements.  The destination is given per 8-element
// (16-byte) chunk by `dst_of(chunk) -> T*`, so the same code fills a cache row in global memory and a swizzled row of a
// shared-memory tile.  Called by `nthreads` consecutive threads (tid in [0,nthreads)).
template <typename T, typename DstFn>
__device__ __forceinline__ void rotate_row_to(const T* src, DstFn dst_of, int d, int rotary_dim, bool interleaved,
                                              const T* cosr, const T* sinr, int tid, int nthreads) {
  const int nchunks = d / 8, rot_chunks = rotary_dim / 8;
  if (interleaved || rotary_dim == 0) {
    for (int c = tid; c < nchunks; c += nthreads) {
      uint4 u = *reinterpret_cast<const uint4*>(src + c * 8);
      if (c < rot_chunks) {
        float x[8], cs[8], sn[8];   // only the first 4 of cs / sn are used
        unpack8<T>(u, x);
        // pairs (2j, 2j+1) use cos/sin[c*4 + j]  (rotary.h:19-78)
        uint2 cu = *reinterpret_cast<const uint2*>(cosr + c * 4), su = *reinterpret_cast<const uint2*>(sinr + c * 4);
        uint4 c4 = make_uint4(cu.x, cu.y, 0, 0), s4 = make_uint4(su.x, su.y, 0, 0);
        unpack8<T>(c4, cs); unpack8<T>(s4, sn);
#pragma unroll
        for (int j = 0; j < 4; j++) {
          float a = x[2 * j], b = x[2 * j + 1];
          x[2 * j] = a * cs[j] - b * sn[j];
          x[2 * j + 1] = a * sn[j] + b * cs[j];
        }
        u = pack8<T>(x);
      }
      *reinterpret_cast<uint4*>(dst_of(c)) = u;
    }
  } else {
    // contiguous (NeoX): pairs (j, j + rotary_dim/2) use cos/sin[j]  (rotary.h:82-145)
    const int half_chunks = rot_chunks / 2;
    for (int c = tid; c < nchunks - half_chunks; c += nthreads) {
      if (c < half_chunks) {
        uint4 u0 = *reinterpret_cast<const uint4*>(src + c * 8);
        uint4 u1 = *reinterpret_cast<const uint4*>(src + (c + half_chunks) * 8);
        uint4 cu = *reinterpret_cast<const uint4*>(cosr + c * 8), su = *reinterpret_cast<const uint4*>(sinr + c * 8);
        float x0[8], x1[8], cs[8], sn[8];
        unpack8<T>(u0, x0); unpack8<T>(u1, x1); unpack8<T>(cu, cs); unpack8<T>(su, sn);
#pragma unroll
        for (int j = 0; j < 8; j++) {
          float a = x0[j], b = x1[j];
          x0[j] = a * cs[j] - b * sn[j];
          x1[j] = a * sn[j] + b * cs[j];
        }
        *reinterpret_cast<uint4*>(dst_of(c)) = pack8<T>(x0);
        *reinterpret_cast<uint4*>(dst_of(c + half_chunks)) = pack8<T>(x1);
      } else {
        const int cc = c + half_chunks;   // chunks beyond the rotary part
        *reinterpret_cast<uint4*>(dst_of(cc)) = *reinterpret_cast<const uint4*>(src + cc * 8);
      }
    }
  }
}

template <typename T>
__device__ __forceinline__ void rotate_row(const T* src, T* dst, int d, int rotary_dim, bool interleaved,
                                           const T* cosr, const T* sinr, int tid, int nthreads) {
  rotate_row_to<T>(src, [dst](int c) { return dst + c * 8; }, d, rotary_dim, interleaved, cosr, sinr, tid, nthreads);
}

}  // namespace fa
