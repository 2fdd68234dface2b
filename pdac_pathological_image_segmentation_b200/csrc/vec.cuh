// 16-byte bf16x8 helpers for the bandwidth-bound NHWC kernels (one thread = 8 consecutive channels of one pixel).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace pdac {

struct F8 {
  float v[8];
};

__device__ __forceinline__ F8 unpack8(const uint4& u) {
  F8 r;
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    r.v[2 * i] = __uint_as_float(w[i] << 16);
    r.v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
  return r;
}
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ uint4 pack8(const F8& f) {
  uint4 u;
  u.x = pack2(f.v[0], f.v[1]);
  u.y = pack2(f.v[2], f.v[3]);
  u.z = pack2(f.v[4], f.v[5]);
  u.w = pack2(f.v[6], f.v[7]);
  return u;
}
__device__ __forceinline__ uint4 ld16(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }
// streaming variants: data touched once, keep it out of the way of reused lines
__device__ __forceinline__ uint4 ld16_stream(const void* p) { return __ldcs(reinterpret_cast<const uint4*>(p)); }
__device__ __forceinline__ void st16(void* p, const uint4& v) { *reinterpret_cast<uint4*>(p) = v; }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// A pixel-contiguous NHWC view flattened to (pixels, C) with a uniform pixel stride.
struct Flat {
  __nv_bfloat16* ptr;
  long long npix;
  int C;
  long long ps;  // pixel stride in elements
};

}  // namespace pdac
